"""TEST INFRASTRUCTURE ONLY (used by tests/, __graft_entry__.smoke() and bench.py's reference arm): ctypes driver of oracle/_ref/libref_f{F}.so -- the UNMODIFIED reference CUDA rasterizer
built for sm_100a by oracle/build_ref.py (prebuilt here, shipped to the GPU box; /root/reference is not needed at run
time).  Device memory comes from torch.  Never imported by the product package."""
import ctypes as C
import os
from types import SimpleNamespace

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_libs = {}


def lib_path(F):
    return os.path.join(_HERE, "_ref", "libref_f%d.so" % F)


def available(F=3):
    return os.path.exists(lib_path(F))


class RefState(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "depths", "clamped", "internal_radii", "means2D", "cov3D", "conic_opacity", "rgb", "point_offsets",
        "tiles_touched", "ranges", "n_contrib", "accum_alpha", "point_list_keys_unsorted", "point_list_keys",
        "point_list_unsorted", "point_list")]


def _lib(F):
    if F not in _libs:
        L = C.CDLL(lib_path(F))
        L.ref_create.restype = C.c_void_p
        L.ref_destroy.argtypes = [C.c_void_p]
        L.ref_forward.restype = C.c_int
        L.ref_backward.restype = C.c_int
        _libs[F] = L
    return _libs[F]


def _p(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _view(ptr, n, dtype, device):
    """Copy n elements of `dtype` from a raw device pointer into a fresh torch tensor."""
    out = torch.empty(n, dtype=dtype, device=device)
    if n:
        L = next(iter(_libs.values()))
        L.ref_copy_d2d.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        assert L.ref_copy_d2d(out.data_ptr(), ptr, n * out.element_size()) == 0
    return out


class Ref:
    """One reference rasterizer instance (owns the three reference buffers between forward and backward)."""

    def __init__(self, F):
        self.F = F
        self.L = _lib(F)
        assert self.L.ref_lang_channels() == F
        self.h = C.c_void_p(self.L.ref_create())

    def __del__(self):
        try:
            self.L.ref_destroy(self.h)
        except Exception:
            pass

    def forward(self, s, include_feature=True, colors_precomp=None, cov3D_precomp=None, dev="cuda"):
        """s: scene namespace of numpy arrays (4dlangsplat_b200/synth.py).  Returns outputs + intermediates (torch)."""
        t = lambda a: None if a is None else torch.as_tensor(np.ascontiguousarray(a), device=dev)
        self.s = s
        self.inp = i = SimpleNamespace(
            means3D=t(s.means3D), shs=None if colors_precomp is not None else t(s.shs), colors=t(colors_precomp),
            lang=t(s.language_feature), opac=t(s.opacities), scales=None if cov3D_precomp is not None else t(s.scales),
            rots=None if cov3D_precomp is not None else t(s.rotations), cov3D=t(cov3D_precomp), view=t(s.viewmatrix),
            proj=t(s.projmatrix), campos=t(s.campos), bg=t(s.bg))
        P, W, H, F = s.P, s.W, s.H, self.F
        M = s.shs.shape[1]
        o = SimpleNamespace()
        o.color = torch.zeros((3, H, W), device=dev)
        o.lang = torch.zeros((F, H, W), device=dev)
        o.depth = torch.zeros((1, H, W), device=dev)
        o.radii = torch.zeros(P, dtype=torch.int32, device=dev)
        self.include_feature = include_feature
        R = self.L.ref_forward(self.h, P, s.sh_degree, M, _p(i.bg), W, H, _p(i.means3D), _p(i.shs), _p(i.colors),
                               _p(i.lang), _p(i.opac), _p(i.scales), C.c_float(s.scale_modifier), _p(i.rots),
                               _p(i.cov3D), _p(i.view), _p(i.proj), _p(i.campos), C.c_float(s.tanfovx),
                               C.c_float(s.tanfovy), 0, _p(o.color), _p(o.lang), _p(o.depth), _p(o.radii), 1,
                               int(include_feature))
        assert R >= 0, "reference forward failed"
        torch.cuda.synchronize()
        o.R = R
        st = RefState()
        self.L.ref_state(self.h, P, R, W * H, C.byref(st))
        T = ((W + 15) // 16) * ((H + 15) // 16)
        o.depths = _view(st.depths, P, torch.float32, dev)
        o.means2D = _view(st.means2D, 2 * P, torch.float32, dev).view(P, 2)
        o.cov3D = _view(st.cov3D, 6 * P, torch.float32, dev).view(P, 6)
        o.conic_opacity = _view(st.conic_opacity, 4 * P, torch.float32, dev).view(P, 4)
        o.rgb = _view(st.rgb, 3 * P, torch.float32, dev).view(P, 3)
        o.clamped = _view(st.clamped, 3 * P, torch.uint8, dev).view(P, 3)
        o.tiles_touched = _view(st.tiles_touched, P, torch.int32, dev)
        o.point_offsets = _view(st.point_offsets, P, torch.int32, dev)
        o.ranges = _view(st.ranges, 2 * T, torch.int32, dev).view(T, 2)
        o.n_contrib = _view(st.n_contrib, W * H, torch.int32, dev).view(H, W)
        o.final_T = _view(st.accum_alpha, W * H, torch.float32, dev).view(H, W)
        o.keys = _view(st.point_list_keys, R, torch.int64, dev)
        o.point_list = _view(st.point_list, R, torch.int32, dev)
        o.keys_unsorted = _view(st.point_list_keys_unsorted, R, torch.int64, dev)
        o.point_list_unsorted = _view(st.point_list_unsorted, R, torch.int32, dev)
        self.out = o
        return o

    def forward_raw(self, P, D, M, W, H, bg, means3D, shs, lang, opac, scales, rots, view, proj, campos, tanfovx,
                    tanfovy, out_color, out_lang, out_depth, radii, include_feature=True, scale_modifier=1.0):
        """Lean call for timing: torch CUDA tensors in, no intermediate read-back, debug off.  Returns R."""
        return self.L.ref_forward(self.h, P, D, M, _p(bg), W, H, _p(means3D), _p(shs), _p(None), _p(lang), _p(opac),
                                  _p(scales), C.c_float(scale_modifier), _p(rots), _p(None), _p(view), _p(proj),
                                  _p(campos), C.c_float(tanfovx), C.c_float(tanfovy), 0, _p(out_color), _p(out_lang),
                                  _p(out_depth), _p(radii), 0, int(include_feature))

    def backward_raw(self, P, D, M, R, W, H, bg, means3D, shs, lang, scales, rots, view, proj, campos, tanfovx, tanfovy,
                     radii, dcol, dlang, ddepth, g, include_feature=True, scale_modifier=1.0):
        """Lean backward for timing; `g` is a namespace of pre-zeroed gradient tensors (see backward())."""
        self.L.ref_backward(self.h, P, D, M, R, _p(bg), W, H, _p(means3D), _p(shs), _p(None), _p(lang), _p(scales),
                            C.c_float(scale_modifier), _p(rots), _p(None), _p(view), _p(proj), _p(campos),
                            C.c_float(tanfovx), C.c_float(tanfovy), _p(radii), _p(dcol), _p(dlang), _p(ddepth),
                            _p(g.means2D), _p(g.conic), _p(g.opacity), _p(g.colors), _p(g.lang), _p(g.depths),
                            _p(g.means3D), _p(g.cov3D), _p(g.sh), _p(g.scales), _p(g.rots), 0, int(include_feature))

    def backward(self, dcol, dlang, ddepth, dev="cuda"):
        s, i, o, F = self.s, self.inp, self.out, self.F
        P, W, H = s.P, s.W, s.H
        M = s.shs.shape[1]
        t = lambda a: torch.as_tensor(np.ascontiguousarray(a), device=dev)
        dcol, dlang, ddepth = t(dcol), t(dlang), t(ddepth)
        z = lambda *shape: torch.zeros(shape, device=dev)
        g = SimpleNamespace(means2D=z(P, 3), conic=z(P, 4), opacity=z(P, 1), colors=z(P, 3), lang=z(P, F), depths=z(P),
                            means3D=z(P, 3), cov3D=z(P, 6), sh=z(P, M, 3), scales=z(P, 3), rots=z(P, 4))
        rc = self.L.ref_backward(self.h, P, s.sh_degree, M, o.R, _p(i.bg), W, H, _p(i.means3D), _p(i.shs), _p(i.colors),
                                 _p(i.lang), _p(i.scales), C.c_float(s.scale_modifier), _p(i.rots), _p(i.cov3D),
                                 _p(i.view), _p(i.proj), _p(i.campos), C.c_float(s.tanfovx), C.c_float(s.tanfovy),
                                 _p(o.radii), _p(dcol), _p(dlang), _p(ddepth), _p(g.means2D), _p(g.conic),
                                 _p(g.opacity), _p(g.colors), _p(g.lang), _p(g.depths), _p(g.means3D), _p(g.cov3D),
                                 _p(g.sh), _p(g.scales), _p(g.rots), 1, int(self.include_feature))
        assert rc == 0
        torch.cuda.synchronize()
        return g

    def mark_visible(self, means3D, view, proj):
        P = means3D.shape[0]
        out = torch.zeros(P, dtype=torch.bool, device=means3D.device)
        self.L.ref_mark_visible(P, _p(means3D), _p(view), _p(proj), _p(out))
        torch.cuda.synchronize()
        return out
