"""fdls -- B200-native (sm_100a) differentiable Gaussian rasterizer for 4D LangSplat's hot path.

The directory name starts with a digit, so import it with importlib:

    fdls = importlib.import_module("4dlangsplat_b200")
    from_ref = fdls.diff_gaussian_rasterization      # GaussianRasterizationSettings, GaussianRasterizer, ...

or simply `import diff_gaussian_rasterization` / `import simple_knn` with the repository root on sys.path (the
top-level shims re-export these modules under the names the reference imports:
gaussian_renderer/__init__.py:15, scene/gaussian_model.py:22).
"""
import importlib as _il

from . import _lib  # noqa: F401  (ctypes view of include/fdls_abi.h)

_C = _il.import_module(__name__ + "._C")
diff_gaussian_rasterization = _il.import_module(__name__ + ".diff_gaussian_rasterization")
simple_knn = _il.import_module(__name__ + ".simple_knn")
hexplane = _il.import_module(__name__ + ".hexplane")
sweep = _il.import_module(__name__ + ".sweep")
deformation = _il.import_module(__name__ + ".deformation")
train_ops = _il.import_module(__name__ + ".train_ops")
gaussian_renderer = _il.import_module(__name__ + ".gaussian_renderer")

__all__ = ["_lib", "_C", "diff_gaussian_rasterization", "simple_knn", "hexplane", "sweep", "deformation", "train_ops", "gaussian_renderer"]
