"""Host side of csrc/train_ops.cu: the streaming kernels around the rasterizer in one training
iteration of the reference's train.py, behind interfaces that mirror what they replace.

    masked_l1_loss(pred, gt, mask, weight)     l1_loss(pred*mask, gt*mask) (utils/loss_utils.py:20-21, train.py:283-296)
                                               loss AND image gradient in one pass; the gradient is the tensor the
                                               rasterizer's backward consumes
    get_language_feature(cam, dir, level, ...) Camera.get_language_feature (scene/cameras.py:69-118): the per-pixel
                                               segment-id -> feature gather runs on the GPU, maps are cached on device
    Adam                                       torch.optim.Adam for the reference's parameter groups
                                               (scene/gaussian_model.py:301): every tensor of every group in ONE launch;
                                               state keys are torch's, so the densification code that edits
                                               optimizer.state (:467-574) works on it unchanged
    add_densification_stats, prune_rows        scene/gaussian_model.py:746-748 and :487-520
No CPU path: non-CUDA tensors raise.
"""
import ctypes as C
import os
from collections import OrderedDict

import numpy as np
import torch

from . import _lib


def _stream(dev):
    return C.c_void_p(torch._C._cuda_getCurrentRawStream(dev.index if dev.index is not None else torch.cuda.current_device()))


def _cuda(t, what):
    if not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor (this build has no CPU path)" % what)
    return t if (t.dtype is torch.float32 and t.is_contiguous()) else t.float().contiguous()


# ---- masked L1 --------------------------------------------------------------------------------------------------------
class _MaskedL1(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, gt, mask, weight):
        pred, gt = _cuda(pred, "pred"), _cuda(gt, "gt")
        Cn = pred.shape[-3] if pred.dim() >= 3 else 1
        HW = pred.numel() // Cn
        m = None
        if mask is not None:
            m = _cuda(mask.to(torch.float32), "mask")
            if m.numel() != HW:
                raise RuntimeError("mask must have one value per pixel (1,H,W)")
        loss = torch.zeros((), dtype=torch.float32, device=pred.device)
        need = ctx.needs_input_grad[0]
        grad = torch.empty_like(pred) if need else None
        _lib.check(_lib.load().fdls_masked_l1(Cn, HW, pred.data_ptr(), gt.data_ptr(), m.data_ptr() if m is not None else None,
                                              float(weight), grad.data_ptr() if need else None, loss.data_ptr(),
                                              _stream(pred.device)), "fdls_masked_l1")
        ctx.grad = grad
        return loss

    @staticmethod
    def backward(ctx, g):
        grad = ctx.grad
        if grad is None:
            return None, None, None, None
        # the usual case is g == 1 (the loss is the root, or one term of a sum): hand the stored gradient out as is
        return (grad if g is None else grad * g), None, None, None


def masked_l1_loss(pred, gt, mask=None, weight=None):
    """weight * sum |pred*mask - gt*mask|; weight defaults to 1 / pred.numel(), i.e. the reference's
    l1_loss(pred*mask, gt*mask) = mean(|.|).  pred, gt: (C,H,W); mask: (1,H,W) or None.  For a batch, call per view
    with weight = coefficient / (B*C*H*W) and add the results (train.py:277-296 takes the mean over the concatenation)."""
    if weight is None:
        weight = 1.0 / max(1, pred.numel())
    return _MaskedL1.apply(pred, gt, mask, weight)


# ---- GT language-feature maps ------------------------------------------------------------------------------------------
_gt_cache = OrderedDict()  # (path, level) -> (feature (F,H,W), mask (1,H,W)) on device
GT_CACHE_ENTRIES = int(os.environ.get("FDLS_GT_CACHE", "512"))


def language_feature_name(cam, language_feature_dir, split="train", data_type="nerfies"):
    """File stem of a view's maps (scene/cameras.py:70-90); None for dynerf video views."""
    if data_type == "nerfies":
        real_id = cam.colmap_id * 4 + 1 if split == "train" else cam.colmap_id * 4 + 3 if split == "test" else cam.colmap_id + 1
        return os.path.join(language_feature_dir, "%06d" % real_id)
    if data_type == "dynerf":
        if split == "test":
            assert cam.colmap_id < 300
        elif split == "video":
            return None
        return os.path.join(language_feature_dir, "%s-%04d" % (cam.cam_name, cam.colmap_id % 300))
    raise NotImplementedError


def gather_language_feature(seg_level, feature_map, H, W, device="cuda"):
    """(F,H,W) feature image + (1,H,W) bool mask from one level of the segment-id map (H,W) and the (K,F) feature table."""
    seg = torch.as_tensor(np.ascontiguousarray(seg_level)).to(device=device, dtype=torch.int32).contiguous()
    fmap = torch.as_tensor(np.ascontiguousarray(feature_map)).to(device=device, dtype=torch.float32).contiguous()
    if not seg.is_cuda:
        raise RuntimeError("gather_language_feature needs a CUDA device")
    K, Fd = fmap.shape
    out = torch.empty((Fd, H, W), dtype=torch.float32, device=seg.device)
    mask = torch.empty((1, H, W), dtype=torch.bool, device=seg.device)
    _lib.check(_lib.load().fdls_gt_feature_gather(H * W, K, Fd, seg.data_ptr(), fmap.data_ptr(), out.data_ptr(),
                                                  mask.data_ptr(), _stream(seg.device)), "fdls_gt_feature_gather")
    return out, mask


def get_language_feature(cam, language_feature_dir, feature_level, split="train", data_type="nerfies"):
    """Camera.get_language_feature (scene/cameras.py:69-118) -> (point_feature (F,H,W) cuda, mask (1,H,W) cuda bool)."""
    stem = language_feature_name(cam, language_feature_dir, split, data_type)
    if stem is None:
        return None, None
    if feature_level not in (0, 1, 2, 3):
        raise ValueError("feature_level=", feature_level)
    key = (stem, int(feature_level))
    hit = _gt_cache.get(key)
    if hit is not None:
        _gt_cache.move_to_end(key)
        return hit
    seg_map = np.load(stem + "_s.npy")      # (4, H, W) segment ids per level, -1 = none
    feature_map = np.load(stem + "_f.npy")  # (K, F)
    H, W = int(cam.image_height), int(cam.image_width)
    res = gather_language_feature(seg_map[feature_level, :H, :W], feature_map, H, W)
    _gt_cache[key] = res
    while len(_gt_cache) > GT_CACHE_ENTRIES:
        _gt_cache.popitem(last=False)
    return res


# ---- Adam --------------------------------------------------------------------------------------------------------------
class Adam(torch.optim.Optimizer):
    """torch.optim.Adam(params, lr, betas, eps) without weight decay / amsgrad: one kernel launch per step()."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        lib = _lib.load()
        by_beta = {}
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda:
                    raise RuntimeError("fdls Adam has no CPU path")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] += 1
                g = p.grad if (p.grad.dtype is torch.float32 and p.grad.is_contiguous()) else p.grad.float().contiguous()
                by_beta.setdefault((tuple(group["betas"]), p.device), []).append((p, g, st, group["lr"], group["eps"]))
        for ((b1, b2), dev), items in by_beta.items():
            for i in range(0, len(items), 32):
                chunk = items[i:i + 32]
                n = len(chunk)
                arr = lambda f: (C.c_void_p * n)(*[f(x) for x in chunk])
                _lib.check(lib.fdls_adam_step(
                    n, arr(lambda x: x[0].data_ptr()), arr(lambda x: x[1].data_ptr()),
                    arr(lambda x: x[2]["exp_avg"].data_ptr()), arr(lambda x: x[2]["exp_avg_sq"].data_ptr()),
                    (C.c_int64 * n)(*[x[0].numel() for x in chunk]), (C.c_float * n)(*[float(x[3]) for x in chunk]),
                    (C.c_float * n)(*[float(x[4]) for x in chunk]), (C.c_int32 * n)(*[int(x[2]["step"]) for x in chunk]),
                    float(b1), float(b2), _stream(dev)), "fdls_adam_step")
        return loss


# ---- densification -----------------------------------------------------------------------------------------------------
def add_densification_stats(xyz_gradient_accum, denom, viewspace_point_grad, update_filter):
    """GaussianModel.add_densification_stats (scene/gaussian_model.py:746-748), in place on (P,1) accumulators."""
    P = xyz_gradient_accum.shape[0]
    f = update_filter.to(torch.uint8) if update_filter.dtype is not torch.uint8 else update_filter
    _lib.check(_lib.load().fdls_densify_stats(P, _cuda(viewspace_point_grad, "grad").data_ptr(), f.contiguous().data_ptr(),
                                              xyz_gradient_accum.data_ptr(), denom.data_ptr(),
                                              _stream(xyz_gradient_accum.device)), "fdls_densify_stats")


def prune_rows(keep, tensors):
    """[t[keep] for t in tensors] (prune_points / _prune_optimizer, scene/gaussian_model.py:487-520) with ONE scan of
    the mask shared by all arrays and one compaction launch per array."""
    keep8 = keep.to(torch.uint8).contiguous()
    if not keep8.is_cuda:
        raise RuntimeError("prune_rows has no CPU path")
    incl = torch.cumsum(keep8.to(torch.int32), 0, dtype=torch.int32)
    dest = (incl - keep8.to(torch.int32)).contiguous()
    n_keep = int(incl[-1]) if keep8.numel() else 0
    lib, out = _lib.load(), []
    for t in tensors:
        tc = _cuda(t, "tensor")
        width = tc.numel() // max(1, tc.shape[0])
        o = torch.empty((n_keep,) + tuple(tc.shape[1:]), dtype=torch.float32, device=tc.device)
        if n_keep == 0 or tc.numel() == 0:
            out.append(o)
            continue
        _lib.check(lib.fdls_compact_rows(tc.shape[0], max(1, width), keep8.data_ptr(), dest.data_ptr(), tc.data_ptr(),
                                         o.data_ptr(), _stream(tc.device)), "fdls_compact_rows")
        out.append(o)
    return out
