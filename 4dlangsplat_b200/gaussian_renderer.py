"""Host-side mirror of the reference's render() (gaussian_renderer/__init__.py:19-248): same signature, same result
dict, so render.py / train.py can `from gaussian_renderer import render` unchanged.  SURVEY.md 8f N2.

`pc` is the reference's GaussianModel (duck-typed: get_xyz, _opacity, get_features, get_language_feature, _scaling,
_rotation, _deformation, the three activations, active_sh_degree, get_covariance); `viewpoint_camera` its Camera /
MiniCam (FoVx, FoVy, image_height, image_width, world_view_transform, full_proj_transform, camera_center, time), or the
PanopticSports dict {'camera': GaussianRasterizationSettings, 'time': t}.

What changes underneath:
  * the camera's three host tensors go up as ONE pinned block per camera and are cached on the device (the reference
    copies world_view_transform, full_proj_transform and camera_center separately on every call, :56-59);
  * the per-point time column is one cached (P,1) buffer refilled with a scalar fill, not torch.tensor(t).repeat(P,1)
    through the host (:73);
  * in the fine stages under torch.no_grad() the deformation network runs as HexPlane kernel + ONE tcgen05 MLP
    kernel, with scaling / rotation / opacity activations (:191-193) folded into its epilogue;
  * the rasterizer is this repo's (diff_gaussian_rasterization mirror).
"""
import math
import os

import torch

from .diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

_cam_cache = {}   # id(camera) -> (source tensors' versions, view, proj, campos) on device
_time_cache = {}  # (device, P) -> (P,1) buffer


def _camera_on_device(cam, device):
    srcs = (cam.world_view_transform, cam.full_proj_transform, cam.camera_center)
    if all(t.is_cuda for t in srcs):
        return srcs
    key = id(cam)
    ver = tuple((t.data_ptr(), t._version) for t in srcs)
    hit = _cam_cache.get(key)
    if hit is not None and hit[0] == ver:
        return hit[1]
    block = torch.cat([t.reshape(-1).float() for t in srcs]).pin_memory().to(device, non_blocking=True)
    out = (block[:16].view(4, 4), block[16:32].view(4, 4), block[32:35])
    _cam_cache[key] = (ver, out)
    if len(_cam_cache) > 4096:
        _cam_cache.pop(next(iter(_cam_cache)))
    return out


def _time_column(t, P, device):
    buf = _time_cache.get((device, P))
    if buf is None:
        buf = torch.empty((P, 1), dtype=torch.float32, device=device)
        _time_cache[(device, P)] = buf
        if len(_time_cache) > 16:
            _time_cache.pop(next(iter(_time_cache)))
    if torch.is_grad_enabled():  # autograd may keep the column alive for backward: do not recycle it
        return torch.full((P, 1), float(t), dtype=torch.float32, device=device)
    buf.fill_(float(t))
    return buf


def render(viewpoint_camera, pc, pipe, bg_color, opt=None, scaling_modifier=1.0, override_color=None, stage="fine-lang",
           cam_type=None, args=None):
    """Render the scene; returns {"render", "language_feature_image", "viewspace_points", "visibility_filter", "radii",
    "depth", "coff"} like the reference.  bg_color must be on the GPU."""
    means3D = pc.get_xyz
    device, P = means3D.device, means3D.shape[0]
    screenspace_points = torch.zeros_like(means3D, requires_grad=True) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    include_feature = "base" not in stage
    if cam_type != "PanopticSports":
        view, proj, campos = _camera_on_device(viewpoint_camera, device)
        raster_settings = GaussianRasterizationSettings(
            image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
            tanfovx=math.tan(viewpoint_camera.FoVx * 0.5), tanfovy=math.tan(viewpoint_camera.FoVy * 0.5), bg=bg_color,
            scale_modifier=scaling_modifier, viewmatrix=view, projmatrix=proj, sh_degree=pc.active_sh_degree,
            campos=campos, prefiltered=False, debug=pipe.debug, include_feature=include_feature)
        t = viewpoint_camera.time
    else:
        raster_settings, t = viewpoint_camera["camera"], viewpoint_camera["time"]
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)

    opacity, shs = pc._opacity, pc.get_features
    F_lang = int(os.getenv("language_feature_hiddendim", 3))
    if include_feature:
        lang = pc.get_language_feature
        if os.getenv("nonormalized", "f") == "f":
            lang = lang / (lang.norm(dim=-1, keepdim=True) + 1e-9)
    else:
        lang = torch.zeros((P, F_lang), dtype=opacity.dtype, device=device)
    scales = rotations = cov3D_precomp = None
    if pipe.compute_cov3D_python:
        cov3D_precomp = pc.get_covariance(scaling_modifier)
    else:
        scales, rotations = pc._scaling, pc._rotation

    coff, activated = None, False
    if "coarse" in stage:
        means3D_f, scales_f, rot_f, opacity_f, shs_f, lang_f = means3D, scales, rotations, opacity, shs, lang
    elif "fine" in stage:
        net = pc._deformation
        net.deformation_net.args.no_dlang = 1 if "base" in stage else args.no_dlang
        time = _time_column(t, P, device)
        if hasattr(net, "forward_dynamic") and "activations" in net.forward_dynamic.__code__.co_varnames \
                and scales is not None and not torch.is_grad_enabled():
            means3D_f, scales_f, rot_f, opacity_f, shs_f, lang_f, coff = net.forward_dynamic(
                means3D, scales, rotations, opacity, shs, lang, time, activations=True)
            activated = True
        else:
            means3D_f, scales_f, rot_f, opacity_f, shs_f, lang_f, coff = net(means3D, scales, rotations, opacity, shs,
                                                                              lang, time)
    else:
        raise NotImplementedError
    if not activated:
        scales_f = pc.scaling_activation(scales_f) if scales_f is not None else None
        rot_f = pc.rotation_activation(rot_f) if rot_f is not None else None
        opacity_f = pc.opacity_activation(opacity_f)

    colors_precomp = override_color  # convert_SHs_python is a debugging path of the original 3DGS: SH -> RGB in torch
    if override_color is None and pipe.convert_SHs_python:
        raise NotImplementedError("convert_SHs_python: let the rasterizer evaluate the SH colours")
    # note: the reference passes the ACTIVATED `opacity` (sigmoid of the deformed logits) here, :227
    rendered_image, language_feature_image, radii, depth = rasterizer(
        means3D=means3D_f, means2D=screenspace_points, shs=shs_f if colors_precomp is None else None,
        colors_precomp=colors_precomp, language_feature_precomp=lang_f, opacities=opacity_f, scales=scales_f,
        rotations=rot_f, cov3D_precomp=cov3D_precomp)
    if "base" in stage:
        language_feature_image = None
    return {"render": rendered_image, "language_feature_image": language_feature_image,
            "viewspace_points": screenspace_points, "visibility_filter": radii > 0, "radii": radii, "depth": depth,
            "coff": coff}
