"""Writes tests/golden/deform_*.npz by importing the UNMODIFIED reference module scene/deformation.py (torch CPU).

Run in the build container only (it reads /root/reference):  python tests/golden/make_golden_deform.py
Each fixture holds the reference network's state_dict, seeded inputs (canonical attributes + per-point time), the seven
outputs of deform_network.forward, and autograd's gradients of a seeded scalar w.r.t. every parameter and input.
Stubs: `tkinter` (a stray import, deformation.py:5), `utils.graphics_utils` (only needed by apply_rotation, unused
in these configurations), `scene.grid` (DenseGrid, only with empty_voxel); `scene` itself is a stub package so that
scene/__init__.py's dataset imports are skipped -- scene/hexplane.py and scene/deformation.py load from their files.
"""
import importlib.util
import os
import sys
import types
from types import SimpleNamespace

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("FDLS_REFERENCE", "/root/reference")


def args_for(name):
    base = dict(net_width=64, timebase_pe=4, defor_depth=1, posebase_pe=10, scale_rotation_pe=2, opacity_pe=2,
                timenet_width=64, timenet_output=32, bounds=1.6, no_dx=False, no_grid=False, no_ds=False, no_dr=False,
                no_do=True, no_dshs=True, no_dlang=1, empty_voxel=False, grid_pe=0, static_mlp=False, apply_rotation=False,
                kplanes_config=dict(grid_dimensions=2, input_coordinate_dim=4, output_coordinate_dim=32,
                                    resolution=[8, 8, 8, 5]), multires=[1, 2])
    if name == "hyper":  # arguments/hypernerf/default.py shape, small grids
        base.update(net_width=128, kplanes_config=dict(grid_dimensions=2, input_coordinate_dim=4, output_coordinate_dim=16,
                                                       resolution=[8, 8, 8, 10]), multires=[1, 2, 4])
    if name == "allheads":  # every head on (neu3d-style no_do / no_dshs = False) + the language MLP
        base.update(net_width=128, no_do=False, no_dshs=False, no_dlang=0,
                    kplanes_config=dict(grid_dimensions=2, input_coordinate_dim=4, output_coordinate_dim=16,
                                        resolution=[6, 6, 6, 4]), multires=[1, 2])
    return SimpleNamespace(**base)


def load_reference():
    for name in ("tkinter",):
        sys.modules.setdefault(name, types.SimpleNamespace(W=None))
    sys.modules.setdefault("utils", types.ModuleType("utils"))
    gu = types.ModuleType("utils.graphics_utils")
    gu.apply_rotation = gu.batch_quaternion_multiply = None
    sys.modules["utils.graphics_utils"] = gu
    scene = types.ModuleType("scene")
    scene.__path__ = [os.path.join(REF, "scene")]
    sys.modules["scene"] = scene
    grid = types.ModuleType("scene.grid")
    grid.DenseGrid = None
    sys.modules["scene.grid"] = grid
    for mod in ("hexplane", "deformation"):
        spec = importlib.util.spec_from_file_location("scene." + mod, os.path.join(REF, "scene", mod + ".py"))
        m = importlib.util.module_from_spec(spec)
        sys.modules["scene." + mod] = m
        spec.loader.exec_module(m)
    return sys.modules["scene.deformation"]


def main():
    ref = load_reference()
    for name, P in (("default", 300), ("hyper", 513), ("allheads", 257)):
        torch.manual_seed(4321)
        a = args_for(name)
        net = ref.deform_network(a)
        with torch.no_grad():
            for n, p in net.named_parameters():
                if "grid" in n:
                    p.add_(0.3 * torch.randn_like(p))
                elif n.endswith("bias"):
                    p.copy_(0.1 * torch.randn_like(p))
        F_lang = 3
        inp = dict(point=torch.rand(P, 3) * 3.0 - 1.5, scales=torch.randn(P, 3) * 0.5 - 3.0,
                   rotations=torch.nn.functional.normalize(torch.randn(P, 4)), opacity=torch.randn(P, 1),
                   shs=torch.randn(P, 16, 3) * 0.3, lang=torch.randn(P, F_lang), times_sel=torch.rand(P, 1) * 2 - 1)
        for k in ("point", "scales", "rotations", "opacity", "shs", "lang"):
            inp[k].requires_grad_(True)
        out = net(inp["point"], inp["scales"], inp["rotations"], inp["opacity"], inp["shs"], inp["lang"], inp["times_sel"])
        names = ("pts", "scales", "rotations", "opacity", "shs", "lang")
        seeds = [torch.randn_like(o) for o in out[:6]]
        loss = sum((o * s).sum() for o, s in zip(out[:6], seeds))
        loss.backward()
        d = {"P": np.int32(P)}
        for k, v in vars(a).items():
            if isinstance(v, (int, float, bool)):
                d["arg_" + k] = np.float64(v)
        d["arg_resolution"] = np.array(a.kplanes_config["resolution"]); d["arg_C"] = np.int32(a.kplanes_config["output_coordinate_dim"])
        d["arg_multires"] = np.array(a.multires)
        for k, v in net.state_dict().items():
            d["sd_" + k] = v.detach().numpy()
        for k, v in inp.items():
            d["in_" + k] = v.detach().numpy()
            if v.grad is not None:
                d["gin_" + k] = v.grad.numpy()
        for k, o, s in zip(names, out[:6], seeds):
            d["out_" + k] = o.detach().numpy()
            d["seed_" + k] = s.numpy()
        for k, p in net.named_parameters():
            if p.grad is not None:
                d["gp_" + k] = p.grad.numpy()
        path = os.path.join(HERE, "deform_%s.npz" % name)
        np.savez_compressed(path, **d)
        print(path, os.path.getsize(path))


if __name__ == "__main__":
    sys.exit(main())
