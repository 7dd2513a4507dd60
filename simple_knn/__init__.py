"""Drop-in import name of the reference's simple-knn package (scene/gaussian_model.py:22)."""
import importlib as _il
import sys as _sys

_C = _il.import_module("4dlangsplat_b200.simple_knn._C")
_sys.modules[__name__ + "._C"] = _C
