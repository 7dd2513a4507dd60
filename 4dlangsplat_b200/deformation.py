"""Host-side mirror of the reference's deformation network (scene/deformation.py): `deform_network`, `Deformation`,
`poc_fre`, `initialize_weights` -- same constructor arguments, same parameter / buffer names (so a reference
checkpoint's state_dict loads), same forward signatures and return tuples.  SURVEY.md 8f N1.

What runs underneath:
  * the HexPlane feature field is 4dlangsplat_b200.hexplane.HexPlaneField (one fused CUDA kernel per direction);
  * under torch.no_grad() (render.py's sweeps) the trunk + every active head is ONE persistent tcgen05 kernel
    (csrc/deform_mlp.cu, include/fdls_abi.h fdls_deform_forward): fp16 x 2 operand splits give fp32-grade GEMMs on
    the tensor cores, accumulators live in TMEM, activations stay on the SM between layers, and optionally the
    activations gaussian_renderer/__init__.py:191-193 applies afterwards (exp / normalize / sigmoid) are folded in;
  * with grad enabled the layers run as the reference's own op sequence (F.linear -> cuBLAS, plain library GEMMs)
    through autograd, behind the fused HexPlane forward/backward kernels.
Configurations the fused kernel does not cover (no_grid, grid_pe > 1, defor_depth > 1, net_width not in {64, 128},
static_mlp, empty_voxel, apply_rotation) take the op-sequence path as well; static_mlp / empty_voxel / apply_rotation
need reference modules that are not mirrored here and raise NotImplementedError rather than silently differ.
"""
import ctypes as C
import os

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib
from .hexplane import HexPlaneField

_HEAD_SPECS = (  # attribute name, args flag that disables it, output width, activation code of the fused epilogue
    ("pos_deform", "no_dx", 3, _lib.FDLS_DEFORM_ACT_NONE),
    ("scales_deform", "no_ds", 3, _lib.FDLS_DEFORM_ACT_EXP),
    ("rotations_deform", "no_dr", 4, _lib.FDLS_DEFORM_ACT_NORMALIZE),
    ("opacity_deform", "no_do", 1, _lib.FDLS_DEFORM_ACT_SIGMOID),
    ("shs_deform", "no_dshs", 48, _lib.FDLS_DEFORM_ACT_NONE),
)


def _env_int(name, default):
    return int(os.getenv(name, default))


def _head(width, out_dim, in_dim=None):
    """ReLU -> Linear -> ReLU -> Linear, the shape every head of the reference has (deformation.py:59-63)."""
    return nn.Sequential(nn.ReLU(), nn.Linear(width if in_dim is None else in_dim, width), nn.ReLU(),
                         nn.Linear(width, out_dim))


def poc_fre(input_data, poc_buf):
    """Positional encoding [x, sin(x * 2^i), cos(x * 2^i)] (deformation.py:262-268)."""
    emb = (input_data.unsqueeze(-1) * poc_buf).flatten(-2)
    return torch.cat([input_data, emb.sin(), emb.cos()], -1)


def initialize_weights(m):
    if isinstance(m, nn.Linear):  # deformation.py:254-260 (the bias keeps nn.Linear's default init there too)
        nn.init.xavier_uniform_(m.weight, gain=1)
        if m.bias is not None:
            nn.init.xavier_uniform_(m.weight, gain=1)


class Deformation(nn.Module):
    def __init__(self, D=8, W=256, input_ch=27, input_ch_time=9, grid_pe=0, skips=[], args=None):
        super().__init__()
        self.D, self.W, self.input_ch, self.input_ch_time = D, W, input_ch, input_ch_time
        self.skips, self.grid_pe, self.no_grid, self.args = skips, grid_pe, args.no_grid, args
        self.grid = HexPlaneField(args.bounds, args.kplanes_config, args.multires)
        if args.empty_voxel or args.static_mlp:
            raise NotImplementedError("empty_voxel / static_mlp masks (scene/grid.DenseGrid) are not mirrored")
        self.ratio = 0
        self.create_net()
        self._prep = None  # (key, prepared operand buffer, head list) of the fused kernel

    @property
    def get_aabb(self):
        return self.grid.get_aabb

    def set_aabb(self, xyz_max, xyz_min):
        self.grid.set_aabb(xyz_max, xyz_min)

    @property
    def get_empty_ratio(self):
        return self.ratio

    def create_net(self):
        W = self.W
        grid_dim = self.grid.feat_dim * (3 if self.grid_pe != 0 else 1)
        trunk = [nn.Linear(4 if self.no_grid else grid_dim, W)]
        for _ in range(self.D - 1):
            trunk += [nn.ReLU(), nn.Linear(W, W)]
        self.feature_out = nn.Sequential(*trunk)
        for name, _, k, _ in _HEAD_SPECS:
            setattr(self, name, _head(W, k))
        F_lang = _env_int("language_feature_hiddendim", 3)
        self.lang_deform = nn.Sequential(nn.ReLU(), nn.Linear(self.args.timebase_pe * 2 + 1 + F_lang, W), nn.ReLU(),
                                         nn.Linear(W, W), nn.ReLU(), nn.Linear(W, F_lang))
        self.discrete_coff_generator = _head(W, _env_int("centers_num", 3))

    # ---- the reference's op sequence (autograd; also the fallback for configurations the fused kernel lacks) --------
    def query_time(self, rays_pts_emb, scales_emb, rotations_emb, time_feature, time_emb):
        if self.no_grid:
            h = torch.cat([rays_pts_emb[:, :3], time_emb[:, :1]], -1)
        else:
            h = self.grid(rays_pts_emb[:, :3], time_emb[:, :1])
            if self.grid_pe > 1:
                h = poc_fre(h, self.grid_pe)
        return self.feature_out(h)

    def forward(self, rays_pts_emb, scales_emb=None, rotations_emb=None, opacity=None, shs_emb=None, lang_emb=None,
                time_feature=None, time_sel_emb=None, time_pos_emb=None, init_centers=False):
        if time_sel_emb is None:
            return self.forward_static(rays_pts_emb[:, :3])
        return self.forward_dynamic(rays_pts_emb, scales_emb, rotations_emb, opacity, shs_emb, lang_emb, time_feature,
                                    time_sel_emb, time_pos_emb, init_centers)

    def forward_static(self, rays_pts_emb):
        raise NotImplementedError("forward_static needs static_mlp, which the reference's configurations never enable")

    def _language(self, hidden, coff, lang_emb, time_pos_emb, init_centers):
        """deformation.py:149-173; `coff` is the discrete-coefficient head's output when it already exists."""
        F_lang, n_c = _env_int("language_feature_hiddendim", 3), _env_int("centers_num", 3)
        if os.getenv("use_discrete_lang_f", "f") == "t" and not init_centers:
            centers = lang_emb[:, :F_lang * n_c].view(lang_emb.shape[0], n_c, -1)
            centers = centers / torch.norm(centers, dim=-1, keepdim=True)
            if coff is None:
                coff = self.discrete_coff_generator(hidden)
            lang = torch.matmul(coff.unsqueeze(1), centers).squeeze(1)
            return lang / (torch.norm(lang, dim=1, keepdim=True) + 1e-9), coff
        assert not (init_centers and self.args.no_dlang), " Dlang must be enabled when initialized centers"
        if self.args.no_dlang:
            return lang_emb[:, :F_lang], None
        src = hidden if os.getenv("use_tribute_dlang", "f") == "t" else torch.concat((lang_emb, time_pos_emb), dim=1)
        dlang = self.lang_deform(src)
        lang = dlang if os.getenv("no_resnet", "f") == "t" else lang_emb[:, :F_lang] + dlang
        return lang / (lang.norm(dim=-1, keepdim=True) + 1e-9), None

    def forward_dynamic(self, rays_pts_emb, scales_emb, rotations_emb, opacity_emb, shs_emb, lang_emb, time_feature,
                        time_sel_emb, time_pos_emb, init_centers=False, activations=False):
        """Returns (pts, scales, rotations, opacity, shs, lang_feature, coff) like deformation.py:103-175.
        activations=True (an addition used by this repo's render()): scales / rotations / opacity come back already
        passed through exp / normalize / sigmoid (gaussian_renderer/__init__.py:191-193)."""
        if self.args.apply_rotation:
            raise NotImplementedError("apply_rotation (utils.graphics_utils.batch_quaternion_multiply) is not mirrored")
        if self._fusable(rays_pts_emb):
            return self._forward_fused(rays_pts_emb, scales_emb, rotations_emb, opacity_emb, shs_emb, lang_emb,
                                       time_sel_emb, time_pos_emb, init_centers, activations)
        a = self.args
        hidden = self.query_time(rays_pts_emb, scales_emb, rotations_emb, time_feature, time_sel_emb)
        pts = rays_pts_emb[:, :3] if a.no_dx else rays_pts_emb[:, :3] + self.pos_deform(hidden)
        scales = scales_emb[:, :3] if a.no_ds else scales_emb[:, :3] + self.scales_deform(hidden)
        rotations = rotations_emb[:, :4] if a.no_dr else rotations_emb[:, :4] + self.rotations_deform(hidden)
        opacity = opacity_emb[:, :1] if a.no_do else opacity_emb[:, :1] + self.opacity_deform(hidden)
        shs = shs_emb if a.no_dshs else shs_emb + self.shs_deform(hidden).reshape([shs_emb.shape[0], 16, 3])
        lang, coff = self._language(hidden, None, lang_emb, time_pos_emb, init_centers)
        if activations:
            scales, rotations, opacity = torch.exp(scales), F.normalize(rotations), torch.sigmoid(opacity)
        return pts, scales, rotations, opacity, shs, lang, coff

    # ---- fused tcgen05 path ------------------------------------------------------------------------------------------
    def _fusable(self, x):
        return (x.is_cuda and not torch.is_grad_enabled() and not self.no_grid and self.grid_pe <= 1 and self.D <= 1
                and self.W in (64, 128) and self.grid.feat_dim % 16 == 0 and self.grid.feat_dim <= 128
                and os.getenv("FDLS_DEFORM_FUSED", "1") != "0")

    def _active_heads(self, init_centers):
        heads = [(n, k, act) for n, flag, k, act in _HEAD_SPECS if not getattr(self.args, flag)]
        if os.getenv("use_discrete_lang_f", "f") == "t" and not init_centers:
            heads.append(("discrete_coff_generator", _env_int("centers_num", 3), _lib.FDLS_DEFORM_ACT_NONE))
        return heads

    def _fill(self, a, heads, prep):
        a.G, a.W, a.H = self.grid.feat_dim, self.W, len(heads)
        a.w0, a.b0 = self.feature_out[0].weight.data_ptr(), self.feature_out[0].bias.data_ptr()
        for i, (name, k, _) in enumerate(heads):
            seq = getattr(self, name)
            h = a.head[i]
            h.k = k
            h.w1, h.b1 = seq[1].weight.data_ptr(), seq[1].bias.data_ptr()
            h.w2, h.b2 = seq[3].weight.data_ptr(), seq[3].bias.data_ptr()
        if prep is not None:
            a.prep, a.prep_bytes = prep.data_ptr(), prep.numel()

    def _prepared(self, heads, dev, stream):
        """fp16 x 2 operand chunks of every weight matrix, rebuilt when a parameter changed (version counters)."""
        mats = [self.feature_out[0].weight] + [w for n, _, _ in heads for w in (getattr(self, n)[1].weight,
                                                                               getattr(self, n)[3].weight)]
        key = tuple((m.data_ptr(), m._version) for m in mats) + (tuple(n for n, _, _ in heads),)
        if self._prep is None or self._prep[0] != key:
            lib = _lib.load()
            a = _lib.FdlsDeformArgs()
            self._fill(a, heads, None)
            nbytes = lib.fdls_deform_prep_bytes(C.byref(a))
            if nbytes == 0:
                raise RuntimeError("fdls_deform: unsupported network shape")
            prep = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            a.prep, a.prep_bytes = prep.data_ptr(), nbytes
            _lib.check(lib.fdls_deform_prepare(C.byref(a), stream), "fdls_deform_prepare")
            self._prep = (key, prep)
        return self._prep[1]

    def _forward_fused(self, pts_emb, scales_emb, rot_emb, opacity_emb, shs_emb, lang_emb, time_sel, time_pos_emb,
                       init_centers, activations):
        lib = _lib.load()
        dev = pts_emb.device
        stream = C.c_void_p(torch._C._cuda_getCurrentRawStream(dev.index))
        P = pts_emb.shape[0]
        heads = self._active_heads(init_centers)
        bases = {"pos_deform": pts_emb[:, :3], "scales_deform": scales_emb[:, :3], "rotations_deform": rot_emb[:, :4],
                 "opacity_deform": opacity_emb[:, :1], "shs_deform": shs_emb.reshape(P, -1) if shs_emb is not None else None,
                 "discrete_coff_generator": None}
        out = {}
        if heads:
            feat = self.grid(pts_emb[:, :3].contiguous(), time_sel[:, :1].contiguous())
            a = _lib.FdlsDeformArgs()
            self._fill(a, heads, self._prepared(heads, dev, stream))
            a.P, a.feat = P, feat.data_ptr()
            keep = [feat]
            for i, (name, k, act) in enumerate(heads):
                base = bases[name]
                if base is not None:
                    base = base.float().contiguous()
                    keep.append(base)
                    a.head[i].base = base.data_ptr()
                o = torch.empty((P, k), dtype=torch.float32, device=dev)
                a.head[i].out, a.head[i].act = o.data_ptr(), (act if activations else _lib.FDLS_DEFORM_ACT_NONE)
                out[name] = o
            _lib.check(lib.fdls_deform_forward(C.byref(a), stream), "fdls_deform_forward")
        act = (lambda f, t: f(t)) if activations else (lambda f, t: t)
        pts = out.get("pos_deform", pts_emb[:, :3])
        scales = out["scales_deform"] if "scales_deform" in out else act(torch.exp, scales_emb[:, :3])
        rotations = out["rotations_deform"] if "rotations_deform" in out else act(F.normalize, rot_emb[:, :4])
        opacity = out["opacity_deform"] if "opacity_deform" in out else act(torch.sigmoid, opacity_emb[:, :1])
        shs = out["shs_deform"].view(P, 16, 3) if "shs_deform" in out else shs_emb
        hidden = None
        if not (os.getenv("use_discrete_lang_f", "f") == "t" and not init_centers) and not self.args.no_dlang \
                and os.getenv("use_tribute_dlang", "f") == "t":
            hidden = self.query_time(pts_emb, scales_emb, rot_emb, None, time_sel)  # rare: lang_deform(hidden)
        lang, coff = self._language(hidden, out.get("discrete_coff_generator"), lang_emb, time_pos_emb, init_centers)
        return pts, scales, rotations, opacity, shs, lang, coff

    def get_mlp_parameters(self):
        return [p for n, p in self.named_parameters() if "grid" not in n]

    def get_grid_parameters(self):
        return [p for n, p in self.named_parameters() if "grid" in n]


class deform_network(nn.Module):
    def __init__(self, args):
        super().__init__()
        times_ch = 2 * args.timebase_pe + 1
        self.timenet = nn.Sequential(nn.Linear(times_ch, args.timenet_width), nn.ReLU(),
                                     nn.Linear(args.timenet_width, args.timenet_output))
        self.deformation_net = Deformation(W=args.net_width, D=args.defor_depth,
                                           input_ch=3 + 3 * args.posebase_pe * 2, grid_pe=args.grid_pe,
                                           input_ch_time=args.timenet_output, args=args)
        for name, n in (("time_poc", args.timebase_pe), ("pos_poc", args.posebase_pe),
                        ("rotation_scaling_poc", args.scale_rotation_pe), ("opacity_poc", args.opacity_pe)):
            self.register_buffer(name, torch.FloatTensor([2 ** i for i in range(n)]))
        self.apply(initialize_weights)

    def forward(self, point, scales=None, rotations=None, opacity=None, shs=None, lang=None, times_sel=None,
                init_centers=False):
        return self.forward_dynamic(point, scales, rotations, opacity, shs, lang, times_sel, init_centers)

    @property
    def get_aabb(self):
        return self.deformation_net.get_aabb

    @property
    def get_empty_ratio(self):
        return self.deformation_net.get_empty_ratio

    def forward_static(self, points):
        return self.deformation_net(points)

    def forward_dynamic(self, point, scales=None, rotations=None, opacity=None, shs=None, lang=None, times_sel=None,
                        init_centers=False, activations=False):
        net = self.deformation_net
        # The reference encodes point / scales / rotations with poc_fre (deformation.py:232-236), but the network only
        # ever reads the slices [:, :3] / [:, :4] of those encodings (:76,118-147), which ARE the inputs (poc_fre puts
        # the raw values first): the 63 + 15 + 20 sin/cos columns per point are never used, so neither path builds them
        # (values and gradients are identical).  The time encoding feeds lang_deform only.
        times_emb = None if net.args.no_dlang else poc_fre(times_sel, self.time_poc)
        return net.forward_dynamic(point, scales, rotations, opacity, shs, lang, None, times_sel, times_emb,
                                   init_centers, activations)

    def get_mlp_parameters(self):
        return self.deformation_net.get_mlp_parameters() + list(self.timenet.parameters())

    def get_grid_parameters(self):
        return self.deformation_net.get_grid_parameters()
