import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def _has_gpu():
    try:
        import ctypes as C
        import xvector_b200
        lib = xvector_b200.load()
        ctx = C.c_void_p()
        rc = lib.xv_create(0, b'{"network_type":"tdnn","pooling_type":"statistics_pooling","feature_dim":4}', C.byref(ctx))
        if rc == 0:
            lib.xv_destroy(ctx)
        return rc == 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if not any("gpu" in it.keywords for it in items):
        return
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no B200 visible (libxvector has no CPU fallback)")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)
