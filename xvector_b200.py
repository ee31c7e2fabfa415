"""Import shim: the package directory is ``tf-kaldi-speaker_b200/`` (not a valid Python identifier),
so this module loads it under the importable name ``xvector_b200``.

    import xvector_b200
    from xvector_b200.model.trainer import Trainer
"""
import importlib.util
import os
import sys

_here = os.path.dirname(os.path.abspath(__file__))
_pkg_dir = os.path.join(_here, "tf-kaldi-speaker_b200")
_spec = importlib.util.spec_from_file_location(
    "xvector_b200", os.path.join(_pkg_dir, "__init__.py"), submodule_search_locations=[_pkg_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["xvector_b200"] = _mod
_spec.loader.exec_module(_mod)
