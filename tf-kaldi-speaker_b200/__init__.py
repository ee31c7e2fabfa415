"""xvector_b200: B200-native (sm_100a) implementation of tf-kaldi-speaker's x-vector embedding hot path.

Python keeps the reference's signatures (model/tdnn.py, model/pooling.py, model/common.py, model/trainer.py,
egs/voxceleb/v1/nnet/lib/extract.py, misc/utils.py); every array computation goes through libxvector.so
(include/xvector.h) via ctypes.  There is no CPU fallback.
"""
from ._lib import LIB_PATH, XVectorError, load  # noqa: F401
from .engine import Engine, PinnedArray, config_json  # noqa: F401

__version__ = "0.1.0"
