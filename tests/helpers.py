"""Shared helpers of the GPU parity tests: run the product path through its C ABI on a synthetic scene and unpack
the workspaces with the layout the ABI publishes (fdls_workspace_layout)."""
import importlib
from types import SimpleNamespace

import numpy as np
import torch


def pkg():
    return importlib.import_module("4dlangsplat_b200")


def to_dev(s, dev="cuda"):
    t = lambda a: torch.as_tensor(np.ascontiguousarray(a), device=dev)
    return SimpleNamespace(means3D=t(s.means3D), shs=t(s.shs), lang=t(s.language_feature), opac=t(s.opacities),
                           scales=t(s.scales), rots=t(s.rotations), view=t(s.viewmatrix), proj=t(s.projmatrix),
                           campos=t(s.campos), bg=t(s.bg))


def run_forward(s, include_feature=True, colors_precomp=None, cov3D_precomp=None, dev="cuda"):
    """Returns outputs and decoded intermediates of the product forward (via 4dlangsplat_b200._C)."""
    fd = pkg()
    fd._C.emit_keys = True  # also write the reference-format sorted keys, so they can be compared bit for bit
    d = to_dev(s, dev)
    E = torch.Tensor([])
    colors = E if colors_precomp is None else torch.as_tensor(colors_precomp, device=dev)
    cov = E if cov3D_precomp is None else torch.as_tensor(cov3D_precomp, device=dev)
    sh = d.shs if colors_precomp is None else E
    sc = d.scales if cov3D_precomp is None else E
    ro = d.rots if cov3D_precomp is None else E
    R, color, lang, depth, radii, geom, binning, img = fd._C.rasterize_gaussians(
        d.bg, d.means3D, colors, d.lang, d.opac, sc, ro, s.scale_modifier, cov, d.view, d.proj, s.tanfovx, s.tanfovy,
        s.H, s.W, sh, s.sh_degree, d.campos, False, False, include_feature)
    torch.cuda.synchronize()
    o = SimpleNamespace(R=R, color=color, lang=lang, depth=depth, radii=radii, geom=geom, binning=binning, img=img,
                        dev=d, colors=colors, cov=cov, sh=sh, sc=sc, ro=ro)
    P, W, H = s.P, s.W, s.H
    cap = fd._C._dev_state(torch.device(dev))["last_capacity"] if P > 0 else 0
    L = fd._lib.workspace_layout(P, W, H, cap)
    T = ((W + 15) // 16) * ((H + 15) // 16)

    def arr(buf, off, n, dtype):
        nbytes = n * torch.empty(0, dtype=dtype).element_size()
        return buf[off:off + nbytes].view(dtype)

    rec = arr(geom, L.rec, 16 * P, torch.float32).view(P, 16)
    o.rec = rec
    o.means2D = rec[:, 0:2]
    o.conic_opacity = torch.stack([rec[:, 2], rec[:, 3], rec[:, 4], rec[:, 5]], 1)
    o.depths = rec[:, 6]
    o.rgb = rec[:, 7:10]
    o.tiles_touched = arr(geom, L.tiles_touched, P, torch.int32)
    o.point_offsets = arr(geom, L.point_offsets, P, torch.int32)
    o.clamped = arr(geom, L.clamped, P, torch.uint8)
    o.depth_order = arr(geom, L.depth_order, P, torch.int32)
    o.final_T = arr(img, L.final_T, W * H, torch.float32).view(H, W)
    o.n_contrib = arr(img, L.n_contrib, W * H, torch.int32).view(H, W)
    o.ranges = arr(img, L.ranges, 2 * T, torch.int32).view(T, 2)
    o.tile_max = arr(img, L.tile_max_contrib, T, torch.int32)
    if R > 0:
        o.keys = arr(binning, L.keys, R, torch.int64)
        o.point_list = arr(binning, L.vals, R, torch.int32)
    else:
        o.keys = torch.empty(0, dtype=torch.int64, device=dev)
        o.point_list = torch.empty(0, dtype=torch.int32, device=dev)
    return o


def run_backward(s, o, dcol, dlang, ddepth, include_feature=True, dev="cuda"):
    fd = pkg()
    d = o.dev
    t = lambda a: torch.as_tensor(np.ascontiguousarray(a), device=dev)
    outs = fd._C.rasterize_gaussians_backward(
        d.bg, d.means3D, o.radii, o.colors, d.lang, o.sc, o.ro, s.scale_modifier, o.cov, d.view, d.proj, s.tanfovx,
        s.tanfovy, t(dcol), t(dlang) if include_feature else torch.zeros(1, device=dev), t(ddepth), o.sh, s.sh_degree,
        d.campos, o.geom, o.R, o.binning, o.img, False, include_feature)
    torch.cuda.synchronize()
    names = ("means2D", "colors", "lang", "opacity", "means3D", "cov3D", "sh", "scales", "rots")
    return SimpleNamespace(**dict(zip(names, outs)))


def rel_close(a, b, rtol, atol_frac=1e-6):
    """|a-b| <= rtol*|b| + atol_frac*max|b| elementwise; returns (ok, worst relative excess info)."""
    a = a.double().flatten()
    b = b.double().flatten()
    scale = b.abs().max().item() if b.numel() else 0.0
    tol = rtol * b.abs() + atol_frac * scale
    err = (a - b).abs()
    bad = err > tol
    worst = (err / (b.abs() + atol_frac * scale + 1e-300)).max().item() if b.numel() else 0.0
    return (not bool(bad.any())), worst, int(bad.sum())


def grad_close(a, b, rtol=1e-4, floor_frac=1e-5):
    """Gradient comparison that tolerates summation-order noise (the reference accumulates with float atomics in
    arbitrary order, this build with warp-tree sums; ill-conditioned entries -- differences of nearly equal sums in
    the rotation / covariance chain -- amplify that noise):
      (1) >= 99.9% of the elements meet the strict bound |a-b| <= rtol*|b_ij| + floor_frac*||b||_inf;
      (2) EVERY element meets |a-b| <= rtol*max(|b_ij|, ||b_i||_inf) + 5*rtol*||b||_inf (the few ill-conditioned
          entries; test_backward_accuracy_vs_float64 shows they are noise of the reference as much as of this build);
      (3) everything is finite.
    Returns (ok, message)."""
    a = a.double().reshape(a.shape[0], -1) if a.ndim > 1 else a.double().reshape(-1, 1)
    b = b.double().reshape(a.shape)
    gmax = b.abs().max().item() if b.numel() else 0.0
    row = b.abs().max(dim=1, keepdim=True).values
    err = (a - b).abs()
    tol_all = rtol * torch.maximum(b.abs(), row) + 5 * rtol * gmax
    tol_el = rtol * b.abs() + floor_frac * gmax
    n_bad = int((err > tol_all).sum())
    frac_strict = 1.0 - float((err > tol_el).sum()) / max(1, b.numel())
    ok = n_bad == 0 and frac_strict >= 0.999 and bool(torch.isfinite(a).all())
    return ok, "violations=%d, strict elementwise pass fraction=%.5f, max abs err=%.3g (max |ref|=%.3g)" % (
        n_bad, frac_strict, err.max().item() if b.numel() else 0.0, gmax)
