"""GPU tests of SURVEY 8f N2-N4: masked-L1 loss kernel, GT language-feature gather, fused Adam, densification helpers and
the render() glue -- each against the reference's own formulation in stock torch / numpy on the same inputs."""
import importlib
import os
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fdls():
    return importlib.import_module("4dlangsplat_b200")


@pytest.fixture(scope="module")
def ops():
    return importlib.import_module("4dlangsplat_b200.train_ops")


@pytest.mark.parametrize("C,H,W,masked", [(3, 48, 64, False), (3, 37, 53, True), (6, 64, 96, True), (1, 5, 7, True)])
def test_masked_l1_matches_torch(ops, C, H, W, masked):
    """l1_loss(pred*mask, gt*mask) (utils/loss_utils.py:20-21 as used at train.py:283-296): value and gradient."""
    g = torch.Generator().manual_seed(C * 100 + H)
    pred = torch.randn(C, H, W, generator=g).cuda().requires_grad_(True)
    gt = torch.randn(C, H, W, generator=g).cuda()
    gt.view(-1)[::7] = pred.detach().view(-1)[::7]  # exact ties: sign(0) = 0
    mask = (torch.rand(1, H, W, generator=g) > 0.3).cuda() if masked else None
    lam = 0.7
    ref_in = pred.detach().clone().requires_grad_(True)
    ref = lam * torch.abs(ref_in * mask - gt * mask).mean() if masked else lam * torch.abs(ref_in - gt).mean()
    ref.backward()
    loss = ops.masked_l1_loss(pred, gt, mask, weight=lam / pred.numel())
    loss.backward()
    torch.testing.assert_close(loss, ref, rtol=2e-6, atol=1e-7)
    torch.testing.assert_close(pred.grad, ref_in.grad, rtol=0, atol=1e-9)
    # as one term of a larger loss
    pred.grad = None
    (2.0 * ops.masked_l1_loss(pred, gt, mask, weight=lam / pred.numel())).backward()
    torch.testing.assert_close(pred.grad, 2.0 * ref_in.grad, rtol=1e-6, atol=1e-9)


def test_gt_language_feature_loader_matches_reference_indexing(ops, tmp_path):
    """Camera.get_language_feature (scene/cameras.py:69-118) restated in torch CPU, vs the GPU gather + cache."""
    H, W, K, F = 40, 56, 23, 3
    rng = np.random.default_rng(5)
    seg_map = rng.integers(-1, K, size=(4, H, W)).astype(np.float32)  # the reference files hold floats
    feature_map = rng.normal(size=(K, F)).astype(np.float32)
    cam = types.SimpleNamespace(colmap_id=3, image_height=H, image_width=W, cam_name="cam00")
    stem = os.path.join(str(tmp_path), "%06d" % (cam.colmap_id * 4 + 1))
    np.save(stem + "_s.npy", seg_map)
    np.save(stem + "_f.npy", feature_map)
    for level in range(4):
        feat, mask = ops.get_language_feature(cam, str(tmp_path), level, split="train", data_type="nerfies")
        # reference formulation (meshgrid gather, negative ids index from the end)
        sm, fm = torch.from_numpy(seg_map), torch.from_numpy(feature_map)
        y, x = torch.meshgrid(torch.arange(0, H), torch.arange(0, W), indexing="ij")
        seg = sm[:, y.reshape(-1, 1), x.reshape(-1, 1)].squeeze(-1).long()
        ref_mask = (seg != -1)[level:level + 1].reshape(1, H, W)
        ref_feat = fm[seg[level:level + 1]].squeeze(0).reshape(H, W, -1).permute(2, 0, 1)
        assert feat.shape == (F, H, W) and mask.shape == (1, H, W) and mask.dtype is torch.bool
        torch.testing.assert_close(feat.cpu(), ref_feat, rtol=0, atol=0)
        assert torch.equal(mask.cpu(), ref_mask)
        again = ops.get_language_feature(cam, str(tmp_path), level, split="train", data_type="nerfies")
        assert again[0].data_ptr() == feat.data_ptr()  # served from the device cache
    assert ops.language_feature_name(cam, "d", "test") == os.path.join("d", "%06d" % 15)
    assert ops.language_feature_name(cam, "d", "video", "dynerf") is None
    assert ops.language_feature_name(cam, "d", "train", "dynerf") == os.path.join("d", "cam00-0003")


def test_fused_adam_matches_torch_adam(ops):
    """The reference's optimiser setup (scene/gaussian_model.py:290-301): several groups, different learning rates,
    eps = 1e-15; 5 steps, one group without gradient in some steps, a learning-rate change in between."""
    g = torch.Generator().manual_seed(9)
    shapes = [(1000, 3), (1000, 1, 3), (1000, 15, 3), (1000, 1), (1000, 4), (77,)]
    lrs = [1.6e-4, 2.5e-3, 1.25e-4, 5e-2, 1e-3, 1e-2]
    mk = lambda: [torch.nn.Parameter(torch.randn(s, generator=torch.Generator().manual_seed(i)).cuda())
                  for i, s in enumerate(shapes)]
    pa, pb = mk(), mk()
    oa = torch.optim.Adam([{"params": [p], "lr": lr, "name": str(i)} for i, (p, lr) in enumerate(zip(pa, lrs))],
                          lr=0.0, eps=1e-15)
    ob = ops.Adam([{"params": [p], "lr": lr, "name": str(i)} for i, (p, lr) in enumerate(zip(pb, lrs))], lr=0.0,
                  eps=1e-15)
    for it in range(5):
        for i, (a, b) in enumerate(zip(pa, pb)):
            if it == 2 and i == 3:
                a.grad = b.grad = None
                continue
            gr = torch.randn(a.shape, generator=g).cuda() * (10.0 ** (i - 3))
            a.grad, b.grad = gr.clone(), gr.clone()
        if it == 3:
            oa.param_groups[0]["lr"] = ob.param_groups[0]["lr"] = 3e-5
        oa.step()
        ob.step()
    for a, b in zip(pa, pb):
        torch.testing.assert_close(b, a, rtol=2e-6, atol=1e-8)
        for key in ("exp_avg", "exp_avg_sq"):  # to rounding of the moments' own scale
            ra = oa.state[a][key]
            torch.testing.assert_close(ob.state[b][key], ra, rtol=2e-6, atol=2e-7 * float(ra.abs().max()))
        assert int(ob.state[b]["step"]) == int(oa.state[a]["step"])


def test_densification_helpers(ops):
    g = torch.Generator().manual_seed(4)
    P = 5000
    vgrad = torch.randn(P, 3, generator=g).cuda()
    filt = (torch.rand(P, generator=g) > 0.4).cuda()
    accum, denom = torch.rand(P, 1, generator=g).cuda(), torch.rand(P, 1, generator=g).cuda().round()
    ra, rd = accum.clone(), denom.clone()
    ra[filt] += torch.norm(vgrad[filt, :2], dim=-1, keepdim=True)  # scene/gaussian_model.py:746-748
    rd[filt] += 1
    ops.add_densification_stats(accum, denom, vgrad, filt)
    torch.testing.assert_close(accum, ra, rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(denom, rd, rtol=0, atol=0)
    tensors = [torch.randn(P, 3, generator=g).cuda(), torch.randn(P, 15, 3, generator=g).cuda(),
               torch.randn(P, 1, generator=g).cuda(), torch.randn(P, generator=g).cuda()]
    out = ops.prune_rows(filt, tensors)
    for o, t in zip(out, tensors):
        assert torch.equal(o, t[filt])
    none = ops.prune_rows(torch.zeros(P, dtype=torch.bool, device="cuda"), tensors[:1])
    assert none[0].shape == (0, 3)


def test_render_glue_matches_reference_composition(fdls):
    """render() (gaussian_renderer/__init__.py:19-248 mirror) on a stand-in GaussianModel: the fine stage under no_grad
    (fused deformation with folded activations) equals deformation -> torch activations -> rasterizer composed by hand,
    and the coarse stage equals the rasterizer on the canonical attributes."""
    from types import SimpleNamespace
    synth = importlib.import_module("4dlangsplat_b200.synth")
    gr = importlib.import_module("4dlangsplat_b200.gaussian_renderer")
    tg = importlib.import_module("test_gpu_deform")
    net, _ = tg.load_case(fdls, "hyper")
    s = synth.scene("small", view_index=1, n_views=4)
    t = lambda a: torch.tensor(a, device="cuda")
    xyz = t(s.means3D)
    net.deformation_net.set_aabb((xyz.max(0).values + 0.1).tolist(), (xyz.min(0).values - 0.1).tolist())
    with torch.no_grad():  # small deltas keep the scene in view
        for n, p in net.named_parameters():
            if n.endswith("_deform.3.weight") or n.endswith("_deform.3.bias"):
                p.mul_(0.02)
    pc = SimpleNamespace(
        get_xyz=xyz, _opacity=torch.logit(t(s.opacities).clamp(1e-4, 1 - 1e-4)), get_features=t(s.shs),
        get_language_feature=t(s.language_feature) * 2.0, _scaling=torch.log(t(s.scales)), _rotation=t(s.rotations) * 1.7,
        _deformation=net, scaling_activation=torch.exp, rotation_activation=torch.nn.functional.normalize,
        opacity_activation=torch.sigmoid, active_sh_degree=s.sh_degree, max_sh_degree=3, get_covariance=None)
    cam = SimpleNamespace(FoVx=2 * np.arctan(s.tanfovx), FoVy=2 * np.arctan(s.tanfovy), image_height=s.H, image_width=s.W,
                          world_view_transform=torch.tensor(s.viewmatrix).view(4, 4),  # host tensors, like Camera
                          full_proj_transform=torch.tensor(s.projmatrix).view(4, 4),
                          camera_center=torch.tensor(s.campos), time=0.37)
    pipe = SimpleNamespace(debug=False, compute_cov3D_python=False, convert_SHs_python=False)
    args = SimpleNamespace(no_dlang=1)
    bg = t(s.bg)
    dgr = fdls.diff_gaussian_rasterization
    rs = dgr.GaussianRasterizationSettings(s.H, s.W, float(np.tan(cam.FoVx * 0.5)), float(np.tan(cam.FoVy * 0.5)), bg, 1.0,
                                           cam.world_view_transform.cuda(), cam.full_proj_transform.cuda(), s.sh_degree,
                                           cam.camera_center.cuda(), False, False, True)
    lang_n = pc.get_language_feature / (pc.get_language_feature.norm(dim=-1, keepdim=True) + 1e-9)
    with torch.no_grad():
        out = gr.render(cam, pc, pipe, bg, None, stage="fine-lang", args=args)
        time = torch.full((xyz.shape[0], 1), cam.time, device="cuda")
        os.environ["FDLS_DEFORM_FUSED"] = "0"
        try:
            m, sc, ro, op, sh, la, _ = net(pc.get_xyz, pc._scaling, pc._rotation, pc._opacity, pc.get_features, lang_n, time)
        finally:
            del os.environ["FDLS_DEFORM_FUSED"]
        want = dgr.GaussianRasterizer(rs)(means3D=m, means2D=None, shs=sh, language_feature_precomp=la,
                                          opacities=torch.sigmoid(op), scales=torch.exp(sc),
                                          rotations=torch.nn.functional.normalize(ro))
        coarse = gr.render(cam, pc, pipe, bg, None, stage="coarse-lang", args=args)
        want_c = dgr.GaussianRasterizer(rs)(means3D=xyz, means2D=None, shs=pc.get_features, language_feature_precomp=lang_n,
                                            opacities=torch.sigmoid(pc._opacity), scales=torch.exp(pc._scaling),
                                            rotations=torch.nn.functional.normalize(pc._rotation))
        base = gr.render(cam, pc, pipe, bg, None, stage="fine-base", args=args)
    assert set(out) == {"render", "language_feature_image", "viewspace_points", "visibility_filter", "radii", "depth", "coff"}
    assert int(out["visibility_filter"].sum()) > 100
    # the fused deformation differs from the op sequence by ~1e-6 in the Gaussians' attributes: images agree to 1e-4
    # except where a pair sits on the alpha >= 1/255 threshold and flips (a handful of pixels, a few 1e-3)
    for a, b in ((out["render"], want[0]), (out["language_feature_image"], want[1])):
        err = (a - b).abs()
        assert float((err > 1e-4).float().mean()) < 1e-3 and float(err.max()) < 2e-2
    assert float((out["radii"] != want[2]).float().mean()) < 1e-3
    for a, b in zip((coarse["render"], coarse["language_feature_image"], coarse["depth"]), (want_c[0], want_c[1], want_c[3])):
        assert torch.equal(a, b)
    assert torch.equal(coarse["radii"], want_c[2])
    assert base["language_feature_image"] is None and base["render"].shape == (3, s.H, s.W)
    # training mode goes through autograd end to end
    pc.get_xyz.requires_grad_(True)
    o = gr.render(cam, pc, pipe, bg, None, stage="fine-lang", args=args)
    o["language_feature_image"].sum().backward()
    assert pc.get_xyz.grad is not None and o["viewspace_points"].grad is not None
