"""Import alias: ``import pyinterp_b200`` loads the package that lives in the
directory ``pangeo-pyinterp_b200/`` (a name Python cannot import directly)."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module("pangeo-pyinterp_b200")
sys.modules[__name__] = _pkg
