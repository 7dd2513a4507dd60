"""Drop-in for the reference's `scene.hexplane` (scene/deformation.py:13 does
`from scene.hexplane import HexPlaneField`): import the same names from here instead."""
import importlib as _il

_impl = _il.import_module("4dlangsplat_b200.hexplane")
HexPlaneField = _impl.HexPlaneField
hexplane_features = _impl.hexplane_features
init_grid_param = _impl.init_grid_param
normalize_aabb = _impl.normalize_aabb
get_normalized_directions = _impl.get_normalized_directions
