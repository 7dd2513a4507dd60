"""Host-side mirror of the reference's `diff_gaussian_rasterization` package
(submodules/4d-langsplat-rasterization/diff_gaussian_rasterization/__init__.py): same public names, argument
meaning, return order, shapes, dtypes and error behaviour, so gaussian_renderer/__init__.py:219-228 runs unchanged.

    GaussianRasterizationSettings   NamedTuple, 13 fields in the reference order          (ref :167-180)
    GaussianRasterizer              nn.Module: forward(...), markVisible(positions)        (ref :182-235)
    rasterize_gaussians             functional entry                                        (ref :21-44)
    _RasterizeGaussians             torch.autograd.Function, 10 inputs -> 4 outputs         (ref :46-165)

Underneath, `_C` is the ctypes binding of the C ABI (include/fdls_abi.h) instead of a pybind module.
"""
from typing import NamedTuple

import torch
from torch import nn

from .. import _C


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    include_feature: bool


def _snapshot(values):
    """Host copies of every tensor argument, for the debug dump (ref cpu_deep_copy_tuple :17-19)."""
    return tuple(v.detach().cpu().clone() if isinstance(v, torch.Tensor) else v for v in values)


def _guarded(fn, args, debug, dump_path, banner):
    """Debug mode reproduces the reference's behaviour: on any exception dump the inputs and re-raise (ref :88-95)."""
    if not debug:
        return fn(*args)
    saved = _snapshot(args)
    try:
        return fn(*args)
    except Exception:
        torch.save(saved, dump_path)
        print(banner)
        raise


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, language_feature_precomp, opacities, scales, rotations,
                cov3Ds_precomp, raster_settings):
        rs = raster_settings
        native_args = (rs.bg, means3D, colors_precomp, language_feature_precomp, opacities, scales, rotations,
                       rs.scale_modifier, cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
                       rs.image_height, rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug,
                       rs.include_feature)
        (num_rendered, color, language_feature, depth, radii, geomBuffer, binningBuffer,
         imgBuffer) = _guarded(_C.rasterize_gaussians, native_args, rs.debug, "snapshot_fw.dump",
                               "\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        # an output the loss never touches arrives as None instead of a zeros tensor (the depth image in every stage of
        # train.py): the backward then skips that output's gradient terms instead of multiplying by zeros
        ctx.set_materialize_grads(False)
        ctx.out_shapes = (color.shape, language_feature.shape)
        ctx.save_for_backward(colors_precomp, language_feature_precomp, means3D, scales, rotations, cov3Ds_precomp,
                              radii, sh, geomBuffer, binningBuffer, imgBuffer)
        # leaves registered with a gradient bucket (dist.GradBucket.register): their gradients are accumulated in place
        ctx.sinks = None
        if _C.grad_sinks:
            sk = tuple(_C.grad_sinks.get(t.data_ptr()) for t in (means2D, colors_precomp, language_feature_precomp, opacities,
                                                           means3D, cov3Ds_precomp, sh, scales, rotations))
            if any(x is not None for x in sk):
                ctx.sinks = sk
        ctx.mark_non_differentiable(radii)
        return color, language_feature, radii, depth

    @staticmethod
    def backward(ctx, grad_out_color, grad_out_language_feature, grad_radii, grad_depth):
        rs = ctx.raster_settings
        (colors_precomp, language_feature_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer,
         binningBuffer, imgBuffer) = ctx.saved_tensors
        if grad_out_color is None and grad_out_language_feature is None and grad_depth is None:
            return (None,) * 10
        if grad_out_color is None:
            grad_out_color = torch.zeros(ctx.out_shapes[0], device=means3D.device)
        if grad_out_language_feature is None:
            grad_out_language_feature = torch.zeros(ctx.out_shapes[1], device=means3D.device)
        if grad_depth is None:
            grad_depth = torch.Tensor([])  # NULL on the native side: "treated as zeros" (include/fdls_abi.h)
        native_args = (rs.bg, means3D, radii, colors_precomp, language_feature_precomp, scales, rotations,
                       rs.scale_modifier, cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
                       grad_out_color, grad_out_language_feature, grad_depth, sh, rs.sh_degree, rs.campos, geomBuffer,
                       ctx.num_rendered, binningBuffer, imgBuffer, rs.debug, rs.include_feature)
        # inputs that do not require grad get None and their kernels' outputs are skipped (the fine-lang stage freezes
        # everything but the language features, scene/gaussian_model.py:258-270); forward's input order is
        # (means3D, means2D, sh, colors, lang, opacities, scales, rotations, cov3D, settings)
        n = ctx.needs_input_grad
        want = (n[1], n[3], n[4], n[5], n[0], n[8], n[2], n[6], n[7])
        out, acc, owners, sh_rgb = None, False, (), None
        if ctx.sinks is not None:
            sk_sh = ctx.sinks[6]
            if sk_sh is not None and sk_sh.view is None:
                # low-rank SH sink (dist.GradBucket(lowrank_sh=...)): this view leaves its masked colour gradient and
                # camera position in a slot; the bucket expands all views' slots into dL_dsh once per iteration
                if want[6]:
                    sh_rgb = sk_sh.owner.take_sh_slot(rs.campos, means3D, rs.sh_degree)
                want = want[:6] + (False,) + want[7:]
            # (means2D, colors, lang, opacities, means3D, cov3D, sh, scales, rotations): the native return order
            out = tuple(k.view if (k is not None and wnt and k.view is not None) else None
                        for k, wnt in zip(ctx.sinks, want))
            owners = {id(k.owner): k.owner for k in ctx.sinks if k is not None}.values()
            acc = not all(b.fresh for b in owners)  # the iteration's first view overwrites, later ones add
            if acc:
                for b in owners:
                    b.zero_if_fresh()
        grads = _guarded(
            lambda *args: _C.rasterize_gaussians_backward(*args, want=want, out=out, accumulate=acc,
                                                          sh_rgb_out=sh_rgb, sh_rgb_header=sh_rgb is not None),
            native_args,
            rs.debug, "snapshot_bw.dump", "\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
        for b in owners:
            b.fresh = False
        if out is not None:  # gradients that went into a sink are not handed to autograd
            grads = tuple(None if o is not None else g for g, o in zip(grads, out))
        (grad_means2D, grad_colors_precomp, grad_language_feature_precomp, grad_opacities, grad_means3D,
         grad_cov3Ds_precomp, grad_sh, grad_scales, grad_rotations) = grads
        # input order of forward(); the settings get no gradient (ref :152-163)
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_language_feature_precomp,
                grad_opacities, grad_scales, grad_rotations, grad_cov3Ds_precomp, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, language_feature_precomp, opacities, scales, rotations,
                        cov3Ds_precomp, raster_settings):
    if not torch.is_grad_enabled():
        # inference (render.py runs under torch.no_grad()): same native call, none of the autograd.Function bookkeeping
        rs = raster_settings
        native_args = (rs.bg, means3D, colors_precomp, language_feature_precomp, opacities, scales, rotations,
                       rs.scale_modifier, cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
                       rs.image_height, rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug,
                       rs.include_feature)
        out = _guarded(_C.rasterize_gaussians, native_args, rs.debug, "snapshot_fw.dump",
                       "\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
        return out[1], out[2], out[4], out[3]
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, language_feature_precomp, opacities,
                                     scales, rotations, cov3Ds_precomp, raster_settings)


def _or_empty(t):
    # the reference stands in a CPU torch.Tensor([]) for every absent optional (ref :208-221); the native side
    # recognises it by data_ptr()==nullptr
    return _EMPTY if t is None else t


_EMPTY = torch.Tensor([])  # never written to: one shared placeholder instead of a new tensor per absent argument


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, language_feature_precomp=None,
                scales=None, rotations=None, cov3D_precomp=None):
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        have_sr = scales is not None and rotations is not None
        any_sr = scales is not None or rotations is not None
        if (not have_sr and cov3D_precomp is None) or (any_sr and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        return rasterize_gaussians(means3D, means2D, _or_empty(shs), _or_empty(colors_precomp),
                                   _or_empty(language_feature_precomp), opacities, _or_empty(scales),
                                   _or_empty(rotations), _or_empty(cov3D_precomp), self.raster_settings)
