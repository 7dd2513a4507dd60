"""Import shim: `from gaussian_renderer import render` (train.py:27, render.py:20 of the reference) resolves to the
B200 build's mirror of render() (4dlangsplat_b200/gaussian_renderer.py).  network_gui (the viewer socket) is outside
the hot path and not provided."""
import importlib as _il

render = _il.import_module("4dlangsplat_b200.gaussian_renderer").render

__all__ = ["render"]
