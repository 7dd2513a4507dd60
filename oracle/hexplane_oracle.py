"""TEST INFRASTRUCTURE ONLY: numpy restatement of the reference's HexPlane feature field.

Only tests/ may import this module.  It follows, function by function:
  normalize_aabb            scene/hexplane.py:20-21
  grid_sample_wrapper       scene/hexplane.py:22-47   (F.grid_sample, bilinear, padding_mode='border',
                                                       align_corners=True; the arithmetic of torch's
                                                       grid_sampler_2d: unnormalise, clip, nw/ne/sw/se)
  interpolate_ms_features   scene/hexplane.py:65-106  (product over the six planes, concat over levels)
  HexPlaneField.get_density scene/hexplane.py:145-165
and their derivatives.  Pinned against tests/golden/hexplane_*.npz, which tests/golden/make_golden_hexplane.py
produced by importing the UNMODIFIED reference module (torch CPU) in the build container.

Planes are in the reference's parameter layout (1, C, H, W); grids[l][ci] for level l and plane ci in
itertools.combinations(range(4), 2) order: xy, xz, xt, yz, yt, zt.
"""
import itertools

import numpy as np

COMBS = list(itertools.combinations(range(4), 2))
f32 = np.float32


def normalize_aabb(pts, aabb):
    pts, aabb = np.asarray(pts, f32), np.asarray(aabb, f32)
    return ((pts - aabb[0]) * (f32(2.0) / (aabb[1] - aabb[0])) - f32(1.0)).astype(f32)


def _axis(coord, size):
    """unnormalise (align_corners=True) + border clip; returns ix, d ix/d coord, floor(ix)."""
    ix = ((coord + f32(1.0)) / f32(2.0)) * f32(size - 1)
    mult = np.full_like(ix, f32(size - 1) / f32(2.0))
    lo, hi = ix <= 0, ix >= f32(size - 1)
    mult[lo | hi] = 0
    ix = np.where(lo, f32(0), np.where(hi, f32(size - 1), ix)).astype(f32)
    return ix, mult, np.floor(ix).astype(np.int64)


def _corners(ix, iy, x0, y0, Wd, Hd):
    fx0, fy0 = x0.astype(f32), y0.astype(f32)
    fx1, fy1 = fx0 + f32(1), fy0 + f32(1)
    w = [(fx1 - ix) * (fy1 - iy), (ix - fx0) * (fy1 - iy), (fx1 - ix) * (iy - fy0), (ix - fx0) * (iy - fy0)]
    xs, ys = [x0, x0 + 1, x0, x0 + 1], [y0, y0, y0 + 1, y0 + 1]
    ok = [(x < Wd) & (y < Hd) for x, y in zip(xs, ys)]
    return w, xs, ys, ok


def sample_plane(grid, u, v):
    """grid (1,C,H,W); u indexes W, v indexes H; returns (n, C) and the pieces the backward needs."""
    g = np.asarray(grid, f32)[0]
    C, Hd, Wd = g.shape
    ix, mx, x0 = _axis(u, Wd)
    iy, my, y0 = _axis(v, Hd)
    w, xs, ys, ok = _corners(ix, iy, x0, y0, Wd, Hd)
    out = np.zeros((u.shape[0], C), f32)
    vals = []
    for k in range(4):
        xv, yv = np.minimum(xs[k], Wd - 1), np.minimum(ys[k], Hd - 1)
        val = g[:, yv, xv].T * ok[k][:, None]  # (n, C), zero where the corner is outside
        vals.append(val.astype(f32))
        out = (out + val * w[k][:, None]).astype(f32)
    return out, (ix, iy, mx, my, x0, y0, w, xs, ys, ok, vals, (C, Hd, Wd))


def forward(pts, time, aabb, grids):
    """HexPlaneField.forward: (P,3), (P,1) -> (P, L*C).  Also returns what backward() re-uses."""
    c = np.concatenate([normalize_aabb(pts, aabb), np.asarray(time, f32).reshape(-1, 1)], axis=1)
    feats, saved = [], []
    for level in grids:
        space, per_plane = f32(1.0), []
        for ci, (a0, a1) in enumerate(COMBS):
            o, s = sample_plane(level[ci], c[:, a0], c[:, a1])
            per_plane.append((o, s))
            space = (space * o).astype(f32)
        feats.append(space)
        saved.append(per_plane)
    return np.concatenate(feats, axis=1), saved


def backward(pts, time, aabb, grids, dL_dout):
    """Gradients w.r.t. every plane (parameter layout), the points and the time column, in float64 accumulation."""
    out, saved = forward(pts, time, aabb, grids)
    P = out.shape[0]
    dcoord = np.zeros((P, 4), np.float64)
    dgrids, o = [], 0
    for level, per_plane in zip(grids, saved):
        C = per_plane[0][0].shape[1]
        g_out = np.asarray(dL_dout, np.float64)[:, o:o + C]
        o += C
        fs = [pp[0].astype(np.float64) for pp in per_plane]
        dlevel = []
        for ci, (a0, a1) in enumerate(COMBS):
            others = np.ones_like(fs[0])
            for cj in range(6):
                if cj != ci:
                    others = others * fs[cj]
            gf = g_out * others  # dL / d interp_ci, (P, C)
            ix, iy, mx, my, x0, y0, w, xs, ys, ok, vals, (Cc, Hd, Wd) = per_plane[ci][1]
            dg = np.zeros((Cc, Hd, Wd), np.float64)
            for k in range(4):
                m = ok[k]
                np.add.at(dg, (slice(None), ys[k][m], xs[k][m]), (gf[m] * w[k][m][:, None].astype(np.float64)).T)
            dlevel.append(dg[None].astype(f32))
            d = [(v.astype(np.float64) * gf).sum(1) for v in vals]  # nw, ne, sw, se
            fx0, fy0 = x0.astype(np.float64), y0.astype(np.float64)
            X, Y = ix.astype(np.float64), iy.astype(np.float64)
            gix = -d[0] * (fy0 + 1 - Y) + d[1] * (fy0 + 1 - Y) - d[2] * (Y - fy0) + d[3] * (Y - fy0)
            giy = -d[0] * (fx0 + 1 - X) - d[1] * (X - fx0) + d[2] * (fx0 + 1 - X) + d[3] * (X - fx0)
            dcoord[:, a0] += mx * gix
            dcoord[:, a1] += my * giy
        dgrids.append(dlevel)
    aabb = np.asarray(aabb, f32)
    scale = (f32(2.0) / (aabb[1] - aabb[0])).astype(np.float64)
    return dgrids, (dcoord[:, :3] * scale).astype(f32), dcoord[:, 3:4].astype(f32)
