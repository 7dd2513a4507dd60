"""Writes tests/golden/hexplane_*.npz by importing the UNMODIFIED reference module scene/hexplane.py (torch CPU).

Run in the build container only (it reads /root/reference):  python tests/golden/make_golden_hexplane.py
The fixture holds seeded inputs (points incl. some outside the aabb and exactly on its faces, times incl. values
outside [-1,1]), the reference's planes, its features, and autograd's gradients for a seeded upstream gradient.
"""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("FDLS_REFERENCE", "/root/reference")

CASES = {  # name: (P, planeconfig, multires, bounds)
    "tiny": (777, dict(grid_dimensions=2, input_coordinate_dim=4, output_coordinate_dim=16, resolution=[8, 8, 8, 10]),
             [1, 2], 1.6),
    "c32": (300, dict(grid_dimensions=2, input_coordinate_dim=4, output_coordinate_dim=32, resolution=[4, 6, 5, 7]),
            [1, 2, 4], 1.0),
}


def load_reference():
    spec = importlib.util.spec_from_file_location("ref_hexplane", os.path.join(REF, "scene", "hexplane.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def main():
    ref = load_reference()
    for name, (P, cfg, multires, bounds) in CASES.items():
        torch.manual_seed(1234)
        field = ref.HexPlaneField(bounds, cfg, multires)
        with torch.no_grad():  # time planes are initialised to exactly 1 (hexplane.py:64-65): make them informative
            for level in field.grids:
                for g in level:
                    g.add_(0.3 * torch.randn_like(g))
        field.set_aabb([1.5, 1.2, 1.7], [-1.4, -1.3, -1.1])
        pts = (torch.rand(P, 3) * 3.4 - 1.7)
        pts[:5] = torch.tensor([1.5, 1.2, 1.7])     # exactly on the max face
        pts[5:10] = torch.tensor([-1.4, -1.3, -1.1])  # exactly on the min face
        time = torch.rand(P, 1) * 2.4 - 1.2
        time[10:13] = 1.0
        time[13:16] = -1.0
        pts.requires_grad_(True)
        time.requires_grad_(True)
        out = field(pts, time)
        g_out = torch.randn_like(out)
        out.backward(g_out)
        d = {"pts": pts.detach().numpy(), "time": time.detach().numpy(), "aabb": field.aabb.detach().numpy(),
             "out": out.detach().numpy(), "g_out": g_out.numpy(), "g_pts": pts.grad.numpy(), "g_time": time.grad.numpy(),
             "L": np.int32(len(field.grids)), "C": np.int32(cfg["output_coordinate_dim"])}
        for l, level in enumerate(field.grids):
            for ci, g in enumerate(level):
                d["grid_%d_%d" % (l, ci)] = g.detach().numpy()
                d["g_grid_%d_%d" % (l, ci)] = g.grad.numpy()
        path = os.path.join(HERE, "hexplane_%s.npz" % name)
        np.savez_compressed(path, **d)
        print(path, out.shape, os.path.getsize(path))


if __name__ == "__main__":
    sys.exit(main())
