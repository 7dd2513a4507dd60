"""Drop-in import name of the reference package (gaussian_renderer/__init__.py:15 does
`from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer`)."""
import importlib as _il

_impl = _il.import_module("4dlangsplat_b200.diff_gaussian_rasterization")
GaussianRasterizationSettings = _impl.GaussianRasterizationSettings
GaussianRasterizer = _impl.GaussianRasterizer
rasterize_gaussians = _impl.rasterize_gaussians
_RasterizeGaussians = _impl._RasterizeGaussians
_C = _impl._C
