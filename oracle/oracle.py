"""TEST INFRASTRUCTURE ONLY: numpy/ctypes front-end of the CPU oracle (oracle/raster_oracle.cpp).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs may import this
module.  It restates CudaRasterizer::Rasterizer::forward/backward (reference
submodules/4d-langsplat-rasterization/cuda_rasterizer/rasterizer_impl.cu:198-348, :352-469) on the CPU.
"""
import ctypes
import os
import subprocess
from types import SimpleNamespace

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None


def build(force=False):
    """Compile the C++ restatement with g++ (IEEE fp32, no FMA contraction, OpenMP over tiles/Gaussians)."""
    src = os.path.join(_HERE, "raster_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["g++", "-O2", "-fopenmp", "-fPIC", "-shared", "-ffp-contract=off", "-o", _LIB_PATH, src])
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.orc_higher_msb.restype = ctypes.c_uint32
    return _lib


def _p(a):
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return ctypes.c_void_p(a.ctypes.data)


def _suf(dt):
    return "f32" if dt == np.float32 else "f64"


def _c(a, dt):
    return None if a is None else np.ascontiguousarray(a, dtype=dt)


def _real(dt, v):
    return ctypes.c_float(v) if dt == np.float32 else ctypes.c_double(v)


def forward(means3D, opacities, view, proj, campos, W, H, tanfovx, tanfovy, bg, shs=None, colors_precomp=None,
            language_feature=None, scales=None, rotations=None, cov3D_precomp=None, sh_degree=3, scale_modifier=1.0,
            include_feature=True, dtype=np.float32):
    """Full forward; returns a namespace with every intermediate the reference keeps in its three buffers."""
    L = lib()
    dt = np.dtype(dtype).type
    s = _suf(dt)
    means3D = _c(means3D, dt)
    P = means3D.shape[0]
    M = 0 if shs is None else shs.shape[1]
    F = 0 if language_feature is None else language_feature.shape[1]
    o = SimpleNamespace(P=P, W=W, H=H, F=F, M=M, D=sh_degree, dtype=dt, include_feature=bool(include_feature))
    o.inputs = SimpleNamespace(means3D=means3D, opacities=_c(opacities, dt).reshape(-1), view=_c(view, dt).reshape(-1),
                               proj=_c(proj, dt).reshape(-1), campos=_c(campos, dt).reshape(-1), bg=_c(bg, dt),
                               shs=_c(shs, dt), colors_precomp=_c(colors_precomp, dt),
                               language_feature=_c(language_feature, dt), scales=_c(scales, dt),
                               rotations=_c(rotations, dt), cov3D_precomp=_c(cov3D_precomp, dt),
                               tanfovx=tanfovx, tanfovy=tanfovy, scale_modifier=scale_modifier)
    i = o.inputs
    o.radii = np.zeros(P, np.int32)
    o.means2D = np.zeros((P, 2), dt)
    o.depths = np.zeros(P, dt)
    o.cov3D = np.zeros((P, 6), dt)
    o.rgb = np.zeros((P, 3), dt)
    o.conic_opacity = np.zeros((P, 4), dt)
    o.tiles_touched = np.zeros(P, np.uint32)
    o.clamped = np.zeros((P, 3), np.uint8)
    getattr(L, "orc_preprocess_fwd_" + s)(
        P, sh_degree, M, _p(means3D), _p(i.scales), _real(dt, scale_modifier), _p(i.rotations), _p(i.opacities),
        _p(i.shs), _p(i.cov3D_precomp), _p(i.colors_precomp), _p(i.view), _p(i.proj), _p(i.campos), W, H,
        _real(dt, tanfovx), _real(dt, tanfovy), _p(o.radii), _p(o.means2D), _p(o.depths), _p(o.cov3D), _p(o.rgb),
        _p(o.conic_opacity), _p(o.tiles_touched), _p(o.clamped))
    o.point_offsets = np.cumsum(o.tiles_touched, dtype=np.uint64).astype(np.uint32)
    R = int(o.tiles_touched.sum(dtype=np.uint64))
    o.num_rendered = R
    gx, gy = (W + 15) // 16, (H + 15) // 16
    o.keys = np.zeros(R, np.uint64)
    o.point_list = np.zeros(R, np.uint32)
    o.ranges = np.zeros((gx * gy, 2), np.uint32)
    getattr(L, "orc_binning_" + s)(P, W, H, _p(o.means2D), _p(o.depths), _p(o.radii), _p(o.tiles_touched),
                                   _p(o.keys), _p(o.point_list), _p(o.ranges))
    render(o)
    return o


def render(o, colors=None):
    """forward.cu:261-395 on the state of `o` (optionally with different per-Gaussian colours)."""
    L = lib()
    dt, s, i = o.dtype, _suf(o.dtype), o.inputs
    W, H, F = o.W, o.H, o.F
    feat = colors if colors is not None else (i.colors_precomp if i.colors_precomp is not None else o.rgb)
    o.feat = _c(feat, dt)
    o.final_T = np.zeros((H, W), dt)
    o.n_contrib = np.zeros((H, W), np.uint32)
    o.color = np.zeros((3, H, W), dt)
    o.language = np.zeros((max(F, 1), H, W), dt) if o.include_feature else np.zeros((1,), dt)
    o.depth = np.zeros((1, H, W), dt)
    getattr(L, "orc_render_fwd_" + s)(W, H, F, _p(o.ranges), _p(o.point_list), _p(o.means2D), _p(o.feat),
                                      _p(i.language_feature), _p(o.depths), _p(o.conic_opacity), _p(i.bg),
                                      int(o.include_feature), _p(o.final_T), _p(o.n_contrib), _p(o.color),
                                      _p(o.language), _p(o.depth))
    return o


def backward(o, dL_dcolor, dL_dlanguage=None, dL_ddepth=None):
    """Rasterizer::backward on the state of a previous forward().  Returns the reference's 9 API gradients plus
    the internal ones (conic, depths).  Per-Gaussian pixel sums are accumulated in float64."""
    L = lib()
    dt, s, i = o.dtype, _suf(o.dtype), o.inputs
    P, W, H, F, M = o.P, o.W, o.H, o.F, o.M
    dpix = _c(dL_dcolor, dt)
    dpix_F = _c(dL_dlanguage, dt) if (o.include_feature and dL_dlanguage is not None) else np.zeros((max(F, 1), H, W), dt)
    dpix_D = _c(dL_ddepth, dt) if dL_ddepth is not None else np.zeros((1, H, W), dt)
    g = SimpleNamespace()
    g_mean2D = np.zeros((P, 2), np.float64)
    g_conic = np.zeros((P, 3), np.float64)
    g_opacity = np.zeros(P, np.float64)
    g_color = np.zeros((P, 3), np.float64)
    g_lang = np.zeros((P, max(F, 1)), np.float64)
    g_depth = np.zeros(P, np.float64)
    getattr(L, "orc_render_bwd_" + s)(W, H, F, _p(o.ranges), _p(o.point_list), _p(i.bg), _p(o.means2D),
                                      _p(o.conic_opacity), _p(o.feat), _p(i.language_feature), _p(o.depths),
                                      _p(o.final_T), _p(o.n_contrib), _p(dpix), _p(dpix_F), _p(dpix_D),
                                      int(o.include_feature), _p(g_mean2D), _p(g_conic), _p(g_opacity), _p(g_color),
                                      _p(g_lang), _p(g_depth))
    g.means2D = np.zeros((P, 3), dt)
    g.means2D[:, :2] = g_mean2D
    g.conic = g_conic.astype(dt)
    g.opacities = g_opacity.astype(dt).reshape(P, 1)
    g.colors = g_color.astype(dt)
    g.language_feature = g_lang[:, :F].astype(dt) if o.include_feature else np.zeros((1,), dt)
    g.depths = g_depth.astype(dt)
    g.means3D = np.zeros((P, 3), dt)
    g.cov3D = np.zeros((P, 6), dt)
    g.sh = np.zeros((P, M, 3), dt)
    g.scales = np.zeros((P, 3), dt)
    g.rotations = np.zeros((P, 4), dt)
    cov3D = i.cov3D_precomp if i.cov3D_precomp is not None else o.cov3D
    getattr(L, "orc_cov2d_bwd_" + s)(P, _p(i.means3D), _p(o.radii), _p(cov3D), W, H, _real(dt, i.tanfovx),
                                     _real(dt, i.tanfovy), _p(i.view), _p(g.conic), _p(g.means3D), _p(g.cov3D))
    getattr(L, "orc_preprocess_bwd_" + s)(P, o.D, M, _p(i.means3D), _p(o.radii), _p(i.shs), _p(o.clamped),
                                          _p(i.scales), _p(i.rotations), _real(dt, i.scale_modifier), _p(i.view),
                                          _p(i.proj), _p(i.campos), _p(g.means2D), _p(g.means3D), _p(g.colors),
                                          _p(g.depths), _p(g.cov3D), _p(g.sh), _p(g.scales), _p(g.rotations))
    return g


def mark_visible(means3D, view):
    means3D = _c(means3D, np.float32)
    out = np.zeros(means3D.shape[0], np.uint8)
    lib().orc_mark_visible_f32(means3D.shape[0], _p(means3D), _p(_c(view, np.float32).reshape(-1)), _p(out))
    return out.astype(bool)


def knn_dist2(points):
    """simple_knn.distCUDA2 semantics (source absent from /root/reference; pinned upstream commit f10d713 of
    zrporz/simple-knn; call site scene/gaussian_model.py:203): mean squared distance to the 3 nearest OTHER
    points.  PARITY UNPINNED (no reference source, tests or vectors); exact k-d tree answer."""
    from scipy.spatial import cKDTree
    pts = np.ascontiguousarray(points, np.float32)
    d, _ = cKDTree(pts.astype(np.float64)).query(pts.astype(np.float64), k=4)
    return (d[:, 1:4] ** 2).mean(axis=1).astype(np.float32)
