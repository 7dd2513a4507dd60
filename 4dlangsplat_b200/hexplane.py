"""Host-side mirror of the reference's scene/hexplane.py, backed by the fused sm_100a kernels in csrc/hexplane.cu
through the C ABI (fdls_hexplane_forward / _backward / _pack / _unpack_grad).

Same names, constructor arguments, parameter layout ((1, C, H, W) planes in `self.grids[level][plane]`, so the
reference's checkpoints load with load_state_dict) and return shapes as scene/hexplane.py:10-170.  What changes is
the execution: the reference's 6*L F.grid_sample launches + 6*L products + torch.cat (and the same number of autograd
nodes backwards) become ONE kernel forward and ONE kernel backward.

No CPU path: CUDA tensors only; fails loudly otherwise (the MLP heads of scene/deformation.py are
not part of this step).
"""
import ctypes as C
import itertools
import weakref
from typing import Optional, Sequence

import torch
import torch.nn as nn

from . import _lib
from ._C import _stream

# (data_ptr, shape) -> (version, channel-last copy, weakref to the plane); one entry per plane parameter, refreshed when the parameter's
# version counter moves (optimizer steps are in-place and bump it)
_packed = {}


def get_normalized_directions(directions):
    """scene/hexplane.py:10-16"""
    return (directions + 1.0) / 2.0


def normalize_aabb(pts, aabb):
    """scene/hexplane.py:20-21 (the fused kernel applies the same three roundings itself)"""
    return (pts - aabb[0]) * (2.0 / (aabb[1] - aabb[0])) - 1.0


def init_grid_param(grid_nd: int, in_dim: int, out_dim: int, reso: Sequence[int], a: float = 0.1, b: float = 0.5):
    """scene/hexplane.py:49-69: one (1, out_dim, reso[second], reso[first]) parameter per coordinate pair; planes that
    involve time start at exactly 1, the others uniform in [a, b]."""
    assert in_dim == len(reso), "Resolution must have same number of elements as input-dimension"
    has_time_planes = in_dim == 4
    assert grid_nd <= in_dim
    grid_coefs = nn.ParameterList()
    for coo_comb in itertools.combinations(range(in_dim), grid_nd):
        p = nn.Parameter(torch.empty([1, out_dim] + [reso[cc] for cc in coo_comb[::-1]]))
        if has_time_planes and 3 in coo_comb:
            nn.init.ones_(p)
        else:
            nn.init.uniform_(p, a=a, b=b)
        grid_coefs.append(p)
    return grid_coefs


def _transpose_batch(fn_name, Cc, hw, src, dst, device):
    n = len(hw)
    if n == 0:
        return
    HW = (C.c_int64 * n)(*hw)
    a = (C.c_void_p * n)(*[t.data_ptr() for t in src])
    b = (C.c_void_p * n)(*[t.data_ptr() for t in dst])
    _lib.check(getattr(_lib.load(), fn_name)(n, Cc, HW, a, b, _stream(device)), fn_name)


def _pack_all(planes):
    """Channel-last (H, W, C) copies of (1, C, H, W) planes, cached per parameter version; every stale plane is
    converted in ONE launch (fdls_hexplane_pack_batch)."""
    out, stale = [], []
    for plane in planes:
        if not plane.is_cuda:
            raise RuntimeError("HexPlane planes must be CUDA tensors (this build has no CPU path)")
        src = plane.detach()
        if src.dtype is not torch.float32 or not src.is_contiguous():
            src = src.float().contiguous()
        key = (src.data_ptr(), tuple(src.shape))
        hit = _packed.get(key)
        if hit is not None and hit[2]() is not plane:  # the address was recycled by another tensor
            hit = None
        if hit is not None and hit[0] == plane._version:
            out.append(hit[1])
            continue
        _, Cc, H, W = src.shape
        # With grad enabled an earlier forward may have saved the cached copy for its backward: a changed parameter
        # then gets a FRESH copy instead of an in-place rewrite behind autograd's back (the reference would raise
        # "modified by an inplace operation"; here the old graph keeps differentiating against the old planes).
        reuse = hit is not None and not torch.is_grad_enabled()
        cl = hit[1] if reuse else torch.empty(H, W, Cc, dtype=torch.float32, device=src.device)
        if hit is None:  # drop the entry when the parameter dies (30 MB per plane set otherwise leak)
            weakref.finalize(plane, _packed.pop, key, None)
        _packed[key] = (plane._version, cl, weakref.ref(plane))
        stale.append((src, cl))
        out.append(cl)
    if stale:
        Cc = stale[0][0].shape[1]
        _transpose_batch("fdls_hexplane_pack_batch", Cc, [s.shape[2] * s.shape[3] for s, _ in stale],
                         [s for s, _ in stale], [c for _, c in stale], stale[0][0].device)
    return out


def _args(pts, time, aabb, packed, L, Cc, reso):
    a = _lib.FdlsHexArgs()
    a.P, a.C, a.L = pts.shape[0], Cc, L
    for l in range(L):
        for k in range(4):
            a.reso[l][k] = reso[l][k]
        for ci in range(6):
            a.planes[l][ci] = packed[l * 6 + ci].data_ptr()
    a.pts, a.time, a.aabb = pts.data_ptr(), time.data_ptr(), aabb.data_ptr()
    return a


class _HexPlaneFeatures(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pts, time, aabb, L, *planes):
        Cc = planes[0].shape[1]
        # plane ci of a level is (1, C, reso[a1], reso[a0]) for (a0, a1) in xy, xz, xt, yz, yt, zt
        reso = [(planes[l * 6 + 0].shape[3], planes[l * 6 + 0].shape[2], planes[l * 6 + 1].shape[2],
                 planes[l * 6 + 2].shape[2]) for l in range(L)]
        if any(p.shape[1] != Cc for p in planes):
            raise NotImplementedError("all planes must have the same channel count")
        packed = _pack_all(planes)
        out = torch.empty(pts.shape[0], L * Cc, dtype=torch.float32, device=pts.device)
        a = _args(pts, time, aabb, packed, L, Cc, reso)
        a.out = out.data_ptr()
        _lib.check(_lib.load().fdls_hexplane_forward(C.byref(a), _stream(pts.device)), "fdls_hexplane_forward")
        ctx.save_for_backward(pts, time, aabb, *packed)
        ctx.meta = (L, Cc, reso, [tuple(p.shape) for p in planes])
        return out

    @staticmethod
    def backward(ctx, g_out):
        pts, time, aabb, *packed = ctx.saved_tensors
        L, Cc, reso, shapes = ctx.meta
        need = ctx.needs_input_grad
        g_out = g_out.float().contiguous()
        a = _args(pts, time, aabb, packed, L, Cc, reso)
        a.dL_dout = g_out.data_ptr()
        g_pts = torch.empty_like(pts) if need[0] else None
        g_time = torch.empty_like(time) if need[1] else None
        if g_pts is not None:
            a.dL_dpts = g_pts.data_ptr()
        if g_time is not None:
            a.dL_dtime = g_time.data_ptr()
        g_cl = [None] * len(packed)
        want = [i for i in range(len(packed)) if need[4 + i]]
        if want:  # one zero-fill for all channel-last plane gradients
            flat = torch.zeros(sum(packed[i].numel() for i in want), dtype=torch.float32, device=pts.device)
            o = 0
            for i in want:
                g_cl[i] = flat[o:o + packed[i].numel()]
                o += packed[i].numel()
                a.dL_dplanes[i // 6][i % 6] = g_cl[i].data_ptr()
        st = _stream(pts.device)
        _lib.check(_lib.load().fdls_hexplane_backward(C.byref(a), st), "fdls_hexplane_backward")
        g_planes = [None] * len(packed)
        for i in want:
            g_planes[i] = torch.empty(shapes[i], dtype=torch.float32, device=pts.device)
        _transpose_batch("fdls_hexplane_unpack_grad_batch", Cc, [shapes[i][2] * shapes[i][3] for i in want],
                         [g_cl[i] for i in want], [g_planes[i] for i in want], pts.device)
        return (g_pts, g_time, None, None, *g_planes)


def hexplane_features(pts, timestamps, aabb, ms_grids):
    """Fused HexPlaneField.get_density (scene/hexplane.py:145-165): pts (..., 3), timestamps (..., 1), aabb (2, 3),
    ms_grids = sequence of levels, each a sequence of six (1, C, H, W) planes.  Returns (n, L*C)."""
    if timestamps is None:
        raise NotImplementedError("this build fuses the 4-D (space + time) HexPlane only; timestamps are required")
    for lv in ms_grids:
        if len(lv) != 6:
            raise NotImplementedError("grid_dimensions=2 over input_coordinate_dim=4 (six planes per level) only")
    if not pts.is_cuda:
        raise RuntimeError("pts must be a CUDA tensor (this build has no CPU path)")
    pts = pts.reshape(-1, 3)
    time = timestamps.reshape(-1, 1)
    if pts.dtype is not torch.float32 or not pts.is_contiguous():
        pts = pts.float().contiguous()
    if time.dtype is not torch.float32 or not time.is_contiguous():
        time = time.float().contiguous()
    aabb = aabb.detach().to(device=pts.device, dtype=torch.float32).contiguous()
    planes = [p for lv in ms_grids for p in lv]
    return _HexPlaneFeatures.apply(pts, time, aabb, len(ms_grids), *planes)


class HexPlaneField(nn.Module):
    """scene/hexplane.py:109-170 with the same constructor, attributes and methods."""

    def __init__(self, bounds, planeconfig, multires) -> None:
        super().__init__()
        aabb = torch.tensor([[bounds, bounds, bounds], [-bounds, -bounds, -bounds]], dtype=torch.float32)
        self.aabb = nn.Parameter(aabb, requires_grad=False)
        self.grid_config = [planeconfig]
        self.multiscale_res_multipliers = multires
        self.concat_features = True
        self.grids = nn.ModuleList()
        self.feat_dim = 0
        for res in self.multiscale_res_multipliers:
            config = self.grid_config[0].copy()
            # multi-resolution on the spatial axes only (hexplane.py:131-134)
            config["resolution"] = [r * res for r in config["resolution"][:3]] + config["resolution"][3:]
            gp = init_grid_param(grid_nd=config["grid_dimensions"], in_dim=config["input_coordinate_dim"],
                                 out_dim=config["output_coordinate_dim"], reso=config["resolution"])
            self.feat_dim += gp[-1].shape[1]
            self.grids.append(gp)

    @property
    def get_aabb(self):
        return self.aabb[0], self.aabb[1]

    def set_aabb(self, xyz_max, xyz_min):
        aabb = torch.tensor([xyz_max, xyz_min], dtype=torch.float32, device=self.aabb.device)
        self.aabb = nn.Parameter(aabb, requires_grad=False)

    def get_density(self, pts: torch.Tensor, timestamps: Optional[torch.Tensor] = None):
        return hexplane_features(pts, timestamps, self.aabb, self.grids)

    def forward(self, pts: torch.Tensor, timestamps: Optional[torch.Tensor] = None):
        return self.get_density(pts, timestamps)
