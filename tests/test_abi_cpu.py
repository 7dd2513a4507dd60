"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/fdls_abi.h
declares, host-only entry points behave, and the Python mirror keeps the reference's surface and error behaviour.
No kernel is launched here."""
import ctypes as C
import inspect
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(fdls):
    hdr = open(os.path.join(ROOT, "include", "fdls_abi.h")).read()
    declared = set(re.findall(r"\b(fdls_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 14
    lib = fdls._lib.load()
    for name in declared:
        assert hasattr(lib, name), name
    assert declared == set(fdls._lib.SYMBOLS), declared ^ set(fdls._lib.SYMBOLS)
    assert lib.fdls_abi_version() == 1
    assert lib.fdls_strerror(-2) == b"workspace too small"


def test_struct_layouts_match_header(fdls):
    """Field names and order of the ctypes structs follow the header text."""
    hdr = open(os.path.join(ROOT, "include", "fdls_abi.h")).read()
    for cname, cls in (("FdlsFwdArgs", fdls._lib.FdlsFwdArgs), ("FdlsBwdArgs", fdls._lib.FdlsBwdArgs),
                       ("FdlsLayout", fdls._lib.FdlsLayout), ("FdlsWorkspaceSizes", fdls._lib.FdlsWorkspaceSizes),
                       ("FdlsHexArgs", fdls._lib.FdlsHexArgs), ("FdlsDeformHead", fdls._lib.FdlsDeformHead),
                       ("FdlsDeformArgs", fdls._lib.FdlsDeformArgs)):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (cname, cname), hdr, re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        names = []
        for stmt in body.split(";"):
            stmt = stmt.strip()
            if not stmt:
                continue
            decl = re.sub(r"^(const\s+)?[A-Za-z_0-9]+\s*\**", "", stmt, count=1)
            names += [re.sub(r"\[.*", "", n.strip().lstrip("*").strip()) for n in decl.split(",")]
        assert names == [f[0] for f in cls._fields_], cname


def test_workspace_sizes_are_monotone_and_aligned(fdls):
    a = fdls._lib.workspace_sizes(1000, 64, 48, 5000)
    b = fdls._lib.workspace_sizes(2000, 128, 96, 50000)
    for f in ("geom_bytes", "img_bytes", "binning_bytes"):
        assert getattr(a, f) % 256 == 0 and getattr(b, f) > getattr(a, f) > 0
    L = fdls._lib.workspace_layout(1000, 64, 48, 5000)
    assert L.vals == 0 and L.rec == 0 and L.final_T == 0  # documented anchors
    assert L.keys % 256 == 0 and L.sort_scratch > L.keys_alt
    s = fdls._lib.FdlsWorkspaceSizes()
    assert fdls._lib.load().fdls_workspace_sizes(-1, 64, 48, 0, C.byref(s)) == -1


def test_null_and_bad_arguments_are_rejected_without_touching_the_gpu(fdls):
    lib = fdls._lib.load()
    a = fdls._lib.FdlsFwdArgs()
    assert lib.fdls_forward(C.byref(a), None) == -1  # W,H = 0
    a.P, a.W, a.H, a.F = 0, 16, 16, 7
    assert lib.fdls_forward(C.byref(a), None) == -4  # F > FDLS_MAX_LANG_CHANNELS
    b = fdls._lib.FdlsBwdArgs()
    assert lib.fdls_backward(C.byref(b), None) == -1
    assert lib.fdls_sort_pairs(None, None, None, None, 10, None, 0, 32, None, 0, None) == -1
    assert lib.fdls_knn_dist2(5, None, None, None, 0, None) == -1
    h = fdls._lib.FdlsHexArgs()
    assert lib.fdls_hexplane_forward(C.byref(h), None) == -1  # L = 0
    h.L, h.C, h.P = 2, 12, 0
    assert lib.fdls_hexplane_forward(C.byref(h), None) == -4  # C not in {4, 8, 16, 32}
    h.C, h.P = 16, 5
    assert lib.fdls_hexplane_backward(C.byref(h), None) == -1  # null pointers
    assert lib.fdls_hexplane_pack(16, 10, None, None, None) == -1


def test_python_surface_matches_reference(fdls):
    dgr = fdls.diff_gaussian_rasterization
    assert dgr.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug", "include_feature")
    sig = inspect.signature(dgr.GaussianRasterizer.forward)
    assert list(sig.parameters) == ["self", "means3D", "means2D", "opacities", "shs", "colors_precomp",
                                    "language_feature_precomp", "scales", "rotations", "cov3D_precomp"]
    assert all(p.default is None for n, p in sig.parameters.items() if n not in ("self", "means3D", "means2D", "opacities"))
    assert issubclass(dgr._RasterizeGaussians, torch.autograd.Function)
    assert hasattr(dgr.GaussianRasterizer, "markVisible") and callable(dgr.rasterize_gaussians)
    import diff_gaussian_rasterization as top  # the drop-in import name used by gaussian_renderer/__init__.py:15
    assert top.GaussianRasterizer is dgr.GaussianRasterizer
    from simple_knn._C import distCUDA2  # scene/gaussian_model.py:22
    assert callable(distCUDA2)


def test_argument_validation_messages(fdls):
    dgr = fdls.diff_gaussian_rasterization
    rs = dgr.GaussianRasterizationSettings(16, 16, 0.5, 0.5, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 3,
                                           torch.zeros(3), False, False, True)
    r = dgr.GaussianRasterizer(rs)
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(m, m, torch.zeros(4, 1))
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 16, 3), colors_precomp=m)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 16, 3), scales=m)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(m, m, torch.zeros(4, 1), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4), cov3D_precomp=torch.zeros(4, 6))
    # no CPU fallback: CPU tensors are refused loudly
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        r(m, m, torch.zeros(4, 1), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4),
          language_feature_precomp=torch.zeros(4, 3))


def test_hexplane_surface_matches_reference(fdls):
    """Constructor, attributes and parameter shapes of scene/hexplane.py:109-143 (CPU: construction only)."""
    import hexplane_b200 as top
    cfg = dict(grid_dimensions=2, input_coordinate_dim=4, output_coordinate_dim=16, resolution=[64, 64, 64, 150])
    f = top.HexPlaneField(1.6, cfg, [1, 2, 4])
    assert top.HexPlaneField is fdls.hexplane.HexPlaneField
    assert f.feat_dim == 48 and f.concat_features and len(f.grids) == 3 and all(len(lv) == 6 for lv in f.grids)
    # xy, xz, xt, yz, yt, zt; (1, C, reso[second], reso[first]); only the spatial axes scale with the level
    assert [tuple(p.shape) for p in f.grids[1]] == [(1, 16, 128, 128), (1, 16, 128, 128), (1, 16, 150, 128),
                                                    (1, 16, 128, 128), (1, 16, 150, 128), (1, 16, 150, 128)]
    assert all(bool((f.grids[0][ci] == 1).all()) for ci in (2, 4, 5))  # time planes start at exactly 1
    assert all(0.1 <= float(f.grids[0][ci].detach().min()) and float(f.grids[0][ci].detach().max()) <= 0.5 for ci in (0, 1, 3))
    assert sorted(f.state_dict()) == sorted(["aabb"] + ["grids.%d.%d" % (l, c) for l in range(3) for c in range(6)])
    assert not f.aabb.requires_grad and f.aabb.tolist() == [[1.6000000238418579] * 3, [-1.6000000238418579] * 3]
    f.set_aabb([1.0, 2.0, 3.0], [-1.0, -2.0, -3.0])
    assert [t.tolist() for t in f.get_aabb] == [[1.0, 2.0, 3.0], [-1.0, -2.0, -3.0]]
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        f(torch.zeros(4, 3), torch.zeros(4, 1))
