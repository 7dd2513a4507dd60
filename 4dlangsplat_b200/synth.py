"""Deterministic synthetic scenes of the shapes BASELINE.json names (numpy only).

Camera conventions follow the reference (scene/cameras.py:56-67, utils/graphics_utils.py:38-71):
viewmatrix = getWorld2View2(R, T).T, projmatrix = viewmatrix @ getProjectionMatrix(znear, zfar, fovx, fovy).T,
both stored row-major and read by the kernels in the "transposed" convention (auxiliary.h:58-66).
"""
import math
from types import SimpleNamespace

import numpy as np

CONFIGS = {
    # name: (P, W, H, F)
    "tiny": (2048, 64, 48, 3),
    "small": (20000, 200, 152, 3),
    "cfg1": (100_000, 800, 800, 3),
    "cfg2": (300_000, 536, 960, 3),
    "cfg2_f6": (300_000, 536, 960, 6),
    "cfg4": (600_000, 1352, 1014, 3),
    "cfg5": (3_000_000, 1920, 1080, 6),
}


def _projection(znear, zfar, tanx, tany):
    P = np.zeros((4, 4), np.float64)
    top, right = tany * znear, tanx * znear
    P[0, 0] = 2.0 * znear / (2 * right)
    P[1, 1] = 2.0 * znear / (2 * top)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def _rot(axis, deg):
    a = math.radians(deg)
    c, s = math.cos(a), math.sin(a)
    i, j = [(1, 2), (2, 0), (0, 1)][axis]
    R = np.eye(3)
    R[i, i], R[i, j], R[j, i], R[j, j] = c, -s, s, c
    return R


# Named world->camera poses (R, t): x_cam = R x_world + t.  "rot2" is a large rotation about two axes (plus a roll)
# with a translation, so that no entry of the view matrix is 0 or 1 and the camera centre is far from the origin.
POSES = {
    "rot2": (_rot(0, 25.0) @ _rot(1, -35.0) @ _rot(2, 10.0), np.array([0.7, -0.4, 1.3])),
    "rot3": (_rot(2, -70.0) @ _rot(0, -50.0) @ _rot(1, 115.0), np.array([-2.1, 0.9, -0.6])),
}


def camera(W, H, view_index=0, n_views=1, znear=0.01, zfar=100.0, pose=None):
    """View v: rotate about +y by (v-(B-1)/2)*4 degrees around the point (0,0,6); or a named pose from POSES."""
    tanx = 0.5
    tany = 0.5 * H / W
    if pose is not None:
        Rw2c, t = POSES[pose]
    else:
        ang = math.radians((view_index - (n_views - 1) / 2.0) * 4.0)
        c, s = math.cos(ang), math.sin(ang)
        Rw2c = np.array([[c, 0, -s], [0, 1, 0], [s, 0, c]], np.float64)  # world -> camera rotation
        pivot = np.array([0.0, 0.0, 6.0])
        # x_cam = Rw2c (x - pivot) + pivot
        t = pivot - Rw2c @ pivot
    Rt = np.eye(4)
    Rt[:3, :3] = Rw2c
    Rt[:3, 3] = t
    view = np.float32(Rt).T.astype(np.float32)                       # world_view_transform (transposed storage)
    proj = (view.astype(np.float64) @ _projection(znear, zfar, tanx, tany).T).astype(np.float32)
    campos = np.linalg.inv(view.astype(np.float64))[3, :3].astype(np.float32)
    return SimpleNamespace(W=W, H=H, tanfovx=float(tanx), tanfovy=float(tany), viewmatrix=np.ascontiguousarray(view),
                           projmatrix=np.ascontiguousarray(proj), campos=campos)


def gaussians(P, W, H, F, seed, sh_coeffs=16, pose=None):
    """Positions are drawn in the frame of the identity camera; with a named pose they are moved to the world frame
    of that camera (x_world = R^T (x_cam - t)) so the same frustum stays populated."""
    rng = np.random.default_rng(seed)
    tanx, tany = 0.5, 0.5 * H / W
    z = rng.uniform(2.0, 10.0, P)
    near = rng.random(P) < 0.02
    z[near] = rng.uniform(-1.0, 0.2, near.sum())
    zz = np.where(near, 3.0, z)
    x = zz * tanx * rng.uniform(-1.15, 1.15, P)
    y = zz * tany * rng.uniform(-1.15, 1.15, P)
    means3D = np.stack([x, y, z], 1)
    if pose is not None:
        Rw2c, t = POSES[pose]
        means3D = (means3D - t) @ Rw2c  # row form of R^T (x - t)
    means3D = means3D.astype(np.float32)
    scales = np.exp(rng.normal(math.log(0.02), 0.7, (P, 3)))
    big = rng.random(P) < 0.01
    scales[big] *= 10.0
    q = rng.normal(size=(P, 4))
    rotations = (q / np.linalg.norm(q, axis=1, keepdims=True))
    opacities = 1.0 / (1.0 + np.exp(-rng.normal(0.5, 2.0, (P, 1))))
    shs = np.concatenate([rng.normal(0, 0.5, (P, 1, 3)), rng.normal(0, 0.1, (P, sh_coeffs - 1, 3))], 1)
    lang = rng.normal(size=(P, F))
    lang /= np.linalg.norm(lang, axis=1, keepdims=True)
    f = np.float32
    return SimpleNamespace(P=P, F=F, means3D=means3D, scales=scales.astype(f), rotations=rotations.astype(f),
                           opacities=opacities.astype(f), shs=np.ascontiguousarray(shs.astype(f)),
                           language_feature=np.ascontiguousarray(lang.astype(f)))


def scene(name="tiny", view_index=0, n_views=1, seed=None, P=None, W=None, H=None, F=None, bg=(1.0, 1.0, 1.0),
          pose=None, sh_degree=3, scale_modifier=1.0):
    p0, w0, h0, f0 = CONFIGS[name]
    P, W, H, F = P or p0, W or w0, H or h0, F or f0
    cfg_id = list(CONFIGS).index(name)
    g = gaussians(P, W, H, F, seed if seed is not None else 1000 * cfg_id, pose=pose)
    cam = camera(W, H, view_index, n_views, pose=pose)
    g.__dict__.update(cam.__dict__)
    g.bg = np.asarray(bg, np.float32)
    g.sh_degree = int(sh_degree)  # active degree; the SH block stays (P,16,3) like the reference's max_sh_degree=3
    g.scale_modifier = float(scale_modifier)
    g.name = name
    return g


def upstream_grads(s, n_views=1, seed=7, depth=False):
    """dL/d(color, language, depth) mimicking L1 / masked-L1 losses (train.py:283-296)."""
    rng = np.random.default_rng(seed + 17 * s.P)
    H, W, F = s.H, s.W, s.F
    dcol = (np.sign(rng.normal(size=(3, H, W))) / (3.0 * H * W * n_views)).astype(np.float32)
    mask = rng.random((1, H, W)) < 0.7
    dlang = (np.sign(rng.normal(size=(F, H, W))) / (F * H * W * n_views) * mask).astype(np.float32)
    ddepth = (rng.normal(size=(1, H, W)) / (H * W)).astype(np.float32) if depth else np.zeros((1, H, W), np.float32)
    return dcol, dlang, ddepth
