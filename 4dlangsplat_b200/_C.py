"""Torch-facing binding of the C ABI: the three functions the reference's pybind module `_C` exports
(RAST/ext.cpp:15-19, rasterize_points.cu:35-256), same argument order, same return tuples.

PyTorch is plumbing only here: it owns device memory (outputs, gradients, the three workspaces) and the current
stream; all arithmetic happens in libfdls.so.  There is no CPU or eager fallback: non-CUDA tensors raise.

Instance-count handling (the reference blocks on a D2H copy of num_rendered, rasterizer_impl.cu:285):
  * exact mode: preprocess, wait for R in pinned host memory, size the binning workspace to R exactly.
  * speculative mode (default once a previous frame is known; env FDLS_SPECULATIVE=0 disables): the binning
    workspace is sized from the previous frame's R with head-room, both stages are enqueued back to back, and the host
    only waits for the tiny R copy (which completes right after the scan, while sort/compositing are still running).
    If R overflowed the guess, stage 2 is simply re-run with the right size.  Results are identical in both modes.
"""
import ctypes as C
import functools
import os

import torch

from . import _lib

_state = {}  # per-device: pinned R buffer, last R
# bench.py hooks: ctypes arrays of cudaEvent_t handles (FDLS_FWD_EVENTS / FDLS_BWD_EVENTS entries) or None
profile = {"fwd": None, "bwd": None}
# While a pipelined sweep is active (sweep.render_sweep) this is a list: a no-grad forward then appends its
# "wait for R, re-run stage 2 on overflow" closure here instead of waiting, and returns a LIST whose [0] (num_rendered)
# is -1 until the closure has run.  The sweep runs the closure before it hands the frame's outputs to the caller.
_deferred = None

# Optional high-priority torch stream for stage 1 (preprocess, depth sort, scan) of the next forward; set per frame by
# sweep.render_sweep.  The block scheduler then places the small, latency-bound stage-1 grids of a younger frame ahead
# of the still-pending compositor CTAs of an older one.  Stage 2 stays on the current stream, ordered behind stage 1.
_front = None

# The torch.cuda.Stream object that is current, when the caller knows it (sweep.render_sweep): torch.cuda.current_stream()
# costs ~14 us per call.
_current = None

# parity tests set this: the partition binning then also writes the reference-format sorted keys (not needed to render)
emit_keys = False

# Gradient sinks (dist.GradBucket.register): leaf tensor's data_ptr() -> _Sink.  When an input of the rasterizer is a registered
# leaf, the backward kernels ADD its gradient straight into the sink's tensor (a slice of the rank's all-reduce bucket)
# and autograd receives None for it: no per-view gradient tensors, no autograd accumulation pass, no pack copy.
grad_sinks = {}


class _Sink:
    __slots__ = ("view", "owner")

    def __init__(self, view, owner):
        self.view, self.owner = view, owner  # owner.fresh: nothing has been written to the bucket this iteration


def _dev_state(device):
    key = device.index if device.index is not None else torch.cuda.current_device()
    st = _state.get(key)
    if st is None:
        st = {"last_R": None, "streams": {}}
        _state[key] = st
    return st


def _stream_state(st, stream):
    """Per (device, stream): the pinned {R, overflow} pair the scan kernel stores into, and the event the host waits
    on.  Frames issued on different streams (a render sweep alternating two streams overlaps one frame's
    latency-bound preprocess/sort/scan with the previous frame's compositor) must not share them."""
    ss = st["streams"].get(stream.value)
    if ss is None:
        ss = {"pinned": torch.zeros(2, dtype=torch.int32).pin_memory(), "event": torch.cuda.Event()}
        st["streams"][stream.value] = ss
    return ss


def _ptr(t):
    """Device address or NULL for the reference's empty-tensor placeholders (rasterize_points.cu relies on
    data_ptr()==nullptr for torch.Tensor([]))."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _prep(t, name):
    if t is None or t.numel() == 0:
        return t
    if not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor (this build has no CPU path)" % name)
    if t.dtype is torch.float32 and t.is_contiguous():  # the usual case: no new tensor object
        return t
    return t.float().contiguous()


@functools.lru_cache(maxsize=4096)
def _sizes(P, W, H, capacity):
    """(geom_bytes, img_bytes, binning_bytes) for this shape; cached (one ctypes round trip per new shape)."""
    s = _lib.workspace_sizes(P, W, H, capacity)
    return s.geom_bytes, s.img_bytes, s.binning_bytes


def _round_capacity(n):
    """Capacity guesses are rounded up to 1/16 of the power of two below them (>= 64 Ki instances): at most ~6 % slack,
    and consecutive frames ask the caching allocator for the same few block sizes instead of a new one every frame
    (a fresh size means a cudaMalloc, which stalls the stream for a millisecond at R ~ 20M)."""
    n = int(n)
    g = 1 << max(16, n.bit_length() - 4)
    return (n + g - 1) & ~(g - 1)


def _stream(device=None):
    """Raw cudaStream_t of torch's current stream (the fast private getter; torch.cuda.current_stream() costs ~14 us)."""
    idx = device.index if (device is not None and device.index is not None) else torch.cuda.current_device()
    return C.c_void_p(torch._C._cuda_getCurrentRawStream(idx))


def rasterize_gaussians(background, means3D, colors, language_feature, opacity, scales, rotations, scale_modifier,
                        cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh,
                        degree, campos, prefiltered, debug, include_feature):
    """RasterizeGaussiansCUDA (rasterize_points.cu:35-129).  Returns
    (num_rendered, out_color, out_language_feature, out_depth, radii, geomBuffer, binningBuffer, imgBuffer)."""
    lib = _lib.load()
    if means3D.ndim != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")  # rasterize_points.cu:59-61
    means3D = _prep(means3D, "means3D")
    dev = means3D.device
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    background, colors, language_feature, opacity, scales, rotations, cov3D_precomp, viewmatrix, projmatrix, sh, campos = [
        _prep(t, n) for t, n in ((background, "bg"), (colors, "colors_precomp"), (language_feature, "language_feature"),
                                 (opacity, "opacities"), (scales, "scales"), (rotations, "rotations"),
                                 (cov3D_precomp, "cov3D_precomp"), (viewmatrix, "viewmatrix"),
                                 (projmatrix, "projmatrix"), (sh, "shs"), (campos, "campos"))]
    F = language_feature.size(1) if (language_feature is not None and language_feature.numel() > 0) else 0
    if F > _lib.FDLS_MAX_LANG_CHANNELS:
        raise RuntimeError("language_feature has %d channels; this build supports up to %d" %
                           (F, _lib.FDLS_MAX_LANG_CHANNELS))
    M = sh.size(1) if (sh is not None and sh.numel() > 0) else 0
    f32 = dict(dtype=torch.float32, device=dev)
    out_color = torch.empty((3, H, W), **f32)
    out_depth = torch.empty((1, H, W), **f32)
    # rasterize_points.cu:72-78: a (1,) tensor stands in for the feature image when include_feature is false
    out_lang = torch.empty((F, H, W), **f32) if include_feature else torch.zeros((1,), **f32)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    geom_bytes, img_bytes, _ = _sizes(P, W, H, 0)
    geom = torch.empty(geom_bytes, dtype=torch.uint8, device=dev)
    img = torch.empty(img_bytes, dtype=torch.uint8, device=dev)
    if P == 0:  # rasterize_points.cu:91: kernels skipped, images are the fill values
        out_color.zero_(); out_depth.zero_(); out_lang.zero_()
        return 0, out_color, out_lang, out_depth, radii, geom, torch.empty(0, dtype=torch.uint8, device=dev), img

    st = _dev_state(dev)
    a = _lib.FdlsFwdArgs()
    a.P, a.D, a.M, a.F, a.W, a.H = P, int(degree), M, F, W, H
    a.tanfovx, a.tanfovy, a.scale_modifier = float(tan_fovx), float(tan_fovy), float(scale_modifier)
    a.prefiltered, a.debug, a.include_feature = int(bool(prefiltered)), int(bool(debug)), int(bool(include_feature))
    a.background, a.means3D, a.shs, a.colors_precomp = _ptr(background), _ptr(means3D), _ptr(sh), _ptr(colors)
    a.language_feature, a.opacities, a.scales, a.rotations = _ptr(language_feature), _ptr(opacity), _ptr(scales), _ptr(rotations)
    a.cov3D_precomp, a.viewmatrix, a.projmatrix, a.campos = _ptr(cov3D_precomp), _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos)
    a.out_color, a.out_depth, a.radii = out_color.data_ptr(), out_depth.data_ptr(), radii.data_ptr()
    a.out_language = out_lang.data_ptr() if include_feature else None
    a.geom_ws, a.geom_bytes, a.img_ws, a.img_bytes = geom.data_ptr(), geom_bytes, img.data_ptr(), img_bytes
    stream = _stream(dev)
    ss = _stream_state(st, stream)
    a.num_rendered_host = ss["pinned"].data_ptr()
    a.num_rendered_hint = -1
    a.binning_algo = 1 if os.environ.get("FDLS_BINNING", "partition") == "onesweep" else 0
    a.emit_keys = int(bool(emit_keys or debug))
    if profile["fwd"] is not None:
        a.stage_events = profile["fwd"]

    def run_stage2(capacity, hint):
        bs = _sizes(P, W, H, capacity)[2] if capacity > 0 else 0
        binning = torch.empty(bs, dtype=torch.uint8, device=dev)
        a.binning_ws, a.binning_bytes, a.capacity = (binning.data_ptr() if bs else None), bs, capacity
        a.num_rendered_hint = hint
        _lib.check(lib.fdls_bin_and_render(C.byref(a), stream), "fdls_bin_and_render")
        return binning

    speculative = os.environ.get("FDLS_SPECULATIVE", "1") != "0" and st["last_R"] is not None and not debug
    a.capacity = 0
    result = [None, out_color, out_lang, out_depth, radii, geom, None, img]
    if speculative:
        capacity = _round_capacity(st["last_R"] * 1.25 + 4096)
        a.capacity = capacity
        ev = ss["event"]
        if _front is not None:
            cur = _current if _current is not None else torch.cuda.current_stream(dev)
            _front.wait_stream(cur)  # inputs and recycled workspaces are ordered on the current stream
            _lib.check(lib.fdls_preprocess(C.byref(a), C.c_void_p(_front.cuda_stream)), "fdls_preprocess")
            ev.record(_front)
            cur.wait_event(ev)
        else:
            _lib.check(lib.fdls_preprocess(C.byref(a), stream), "fdls_preprocess")
            ev.record(_current if _current is not None else torch.cuda.current_stream(dev))
        result[6] = run_stage2(capacity, -1)
        defer = _deferred is not None and not torch.is_grad_enabled()
        tstream = (_current if _current is not None else torch.cuda.current_stream(dev)) if defer else None

        def finish():
            ev.synchronize()
            R = int(ss["pinned"][0])
            if R > capacity:  # the guess was too small: bin and composite again into the same outputs
                if tstream is not None:
                    with torch.cuda.stream(tstream):
                        result[6] = run_stage2(R, R)
                else:
                    result[6] = run_stage2(R, R)
            result[0] = R
            st["last_R"] = R
            st["last_capacity"] = int(a.capacity)  # read by the tests to decode the binning workspace

        if defer:  # a pipelined sweep (sweep.render_sweep) checks R later, before it hands the images out
            _deferred.append(finish)
            result[0] = -1
            return result
        finish()
    else:
        _lib.check(lib.fdls_preprocess(C.byref(a), stream), "fdls_preprocess")
        torch.cuda.current_stream(dev).synchronize()
        R = int(ss["pinned"][0])
        result[6] = run_stage2(R, R)
        result[0] = R
        st["last_R"] = R
        st["last_capacity"] = int(a.capacity)
    if debug:  # CHECK_CUDA semantics (auxiliary.h:166-173): surface asynchronous errors here
        torch.cuda.current_stream(dev).synchronize()
    return tuple(result)


def rasterize_gaussians_backward(background, means3D, radii, colors, language_feature, scales, rotations,
                                 scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                                 dL_dout_color, dL_dout_language_feature, dL_dout_depth, sh, degree, campos, geomBuffer,
                                 R, binningBuffer, imageBuffer, debug, include_feature, want=None, out=None,
                                 accumulate=False, sh_rgb_out=None, sh_rgb_header=False):
    """RasterizeGaussiansBackwardCUDA (rasterize_points.cu:131-235).  Returns (dL_dmeans2D, dL_dcolors,
    dL_dlanguage_feature, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations).
    `want` (an addition; None = all, like the reference): 9 booleans in that order -- gradients not wanted are neither
    allocated nor written (the C ABI takes a null pointer for them) and come back as None.
    `out` (an addition): 9 optional tensors in that order to write into instead of allocating; with `accumulate` the
    kernels add to every output (caller-provided ones must hold the running sum, fresh ones are zero-filled here).
    `sh_rgb_out`: optional (P,3) tensor that receives the masked colour gradient (FdlsBwdArgs.dL_dsh_rgb), the
    view-dependent factor of the rank-1 SH gradient; see sh_grad_expand.  `sh_rgb_header`: the kernel also writes the
    slot header {campos, 1.0} into the four floats in front of sh_rgb_out (FdlsBwdArgs.sh_rgb_header)."""
    lib = _lib.load()
    means3D = _prep(means3D, "means3D")
    dev = means3D.device
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)  # rasterize_points.cu:160-161
    colors, language_feature, scales, rotations, cov3D_precomp, sh = [
        _prep(t, n) for t, n in ((colors, "colors_precomp"), (language_feature, "language_feature"),
                                 (scales, "scales"), (rotations, "rotations"), (cov3D_precomp, "cov3D_precomp"),
                                 (sh, "shs"))]
    background, viewmatrix, projmatrix, campos = [
        _prep(t, n) for t, n in ((background, "bg"), (viewmatrix, "viewmatrix"), (projmatrix, "projmatrix"),
                                 (campos, "campos"))]
    dL_dout_color = _prep(dL_dout_color, "dL_dout_color")
    dL_dout_language_feature = _prep(dL_dout_language_feature, "dL_dout_language_feature")
    dL_dout_depth = _prep(dL_dout_depth, "dL_dout_depth")
    F = language_feature.size(1) if (language_feature is not None and language_feature.numel() > 0) else 0
    M = sh.size(1) if (sh is not None and sh.numel() > 0) else 0
    f32 = dict(dtype=torch.float32, device=dev)
    # every element of these is written by the backward kernels (zeros for culled Gaussians): no zero-fill pass
    w = (True,) * 9 if want is None else tuple(bool(x) for x in want)
    o = (None,) * 9 if out is None else out
    mk = torch.zeros if accumulate else torch.empty
    E = lambda i, shape: o[i] if o[i] is not None else mk(shape, **f32)
    g_means2D = E(0, (P, 3)) if w[0] else None
    g_colors = E(1, (P, 3)) if w[1] else None
    g_lang = (E(2, (P, F)) if include_feature else torch.zeros((1,), **f32)) if w[2] else None
    g_opacity = E(3, (P, 1)) if w[3] else None
    g_means3D = E(4, (P, 3)) if w[4] else None
    g_cov3D = E(5, (P, 6)) if w[5] else None
    g_sh = (E(6, (P, M, 3)) if M > 0 else torch.zeros((P, 0, 3), **f32)) if w[6] else None
    g_scales = E(7, (P, 3)) if w[7] else None
    g_rot = E(8, (P, 4)) if w[8] else None
    if P == 0:
        return g_means2D, g_colors, g_lang, g_opacity, g_means3D, g_cov3D, g_sh, g_scales, g_rot
    gb = lib.fdls_grad_workspace_bytes(P)
    grad_ws = torch.empty(gb, dtype=torch.uint8, device=dev)
    a = _lib.FdlsBwdArgs()
    a.P, a.D, a.M, a.F, a.W, a.H = P, int(degree), M, F, W, H
    a.tanfovx, a.tanfovy, a.scale_modifier = float(tan_fovx), float(tan_fovy), float(scale_modifier)
    a.debug, a.include_feature, a.num_rendered = int(bool(debug)), int(bool(include_feature)), int(R)
    a.background, a.means3D, a.shs, a.colors_precomp = _ptr(background), _ptr(means3D), _ptr(sh), _ptr(colors)
    a.language_feature, a.scales, a.rotations, a.cov3D_precomp = _ptr(language_feature), _ptr(scales), _ptr(rotations), _ptr(cov3D_precomp)
    a.viewmatrix, a.projmatrix, a.campos, a.radii = _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos), radii.data_ptr()
    a.geom_ws, a.img_ws, a.binning_ws, a.capacity = geomBuffer.data_ptr(), imageBuffer.data_ptr(), _ptr(binningBuffer), int(R)
    a.dL_dout_color = dL_dout_color.data_ptr()
    a.dL_dout_language = _ptr(dL_dout_language_feature) if include_feature else None
    a.dL_dout_depth = _ptr(dL_dout_depth)
    a.grad_ws, a.grad_bytes = grad_ws.data_ptr(), gb
    dp = lambda t: None if t is None else t.data_ptr()
    a.dL_dmeans3D, a.dL_dmeans2D, a.dL_dcolors = dp(g_means3D), dp(g_means2D), dp(g_colors)
    a.dL_dsh = dp(g_sh) if M > 0 else None
    a.dL_dlanguage = dp(g_lang) if include_feature else None
    a.dL_dopacity, a.dL_dcov3D = dp(g_opacity), dp(g_cov3D)
    a.dL_dscales, a.dL_drotations = dp(g_scales), dp(g_rot)
    a.dL_dconic, a.dL_ddepths = None, None
    a.bwd_algo = 1 if os.environ.get("FDLS_RENDER_BWD", "mma") == "shfl" else 0
    a.accumulate = int(bool(accumulate))
    a.dL_dsh_rgb = dp(sh_rgb_out)
    a.sh_rgb_header = int(bool(sh_rgb_header and sh_rgb_out is not None))
    if profile["bwd"] is not None:
        a.stage_events = profile["bwd"]
    _lib.check(lib.fdls_backward(C.byref(a), _stream(dev)), "fdls_backward")
    if debug:
        torch.cuda.current_stream(dev).synchronize()
    return g_means2D, g_colors, g_lang, g_opacity, g_means3D, g_cov3D, g_sh, g_scales, g_rot


def sh_grad_expand(means3D, slots, degree, out, accumulate=False):
    """out (P,M,3) (+)= sum over the view slots of basis(view direction) (x) dL_dRGB (fdls_sh_grad_expand).
    slots: (B, 4 + 3P) float32: camera position, valid flag, the (P,3) array written through sh_rgb_out."""
    lib = _lib.load()
    P, M = out.size(0), out.size(1)
    assert slots.is_contiguous() and out.is_contiguous() and slots.size(1) >= 4 + 3 * P
    _lib.check(lib.fdls_sh_grad_expand(P, int(degree), M, slots.size(0), _prep(means3D, "means3D").data_ptr(),
                                       slots.data_ptr(), slots.size(1), out.data_ptr(), int(bool(accumulate)),
                                       _stream(out.device)), "fdls_sh_grad_expand")
    return out


def mark_visible(means3D, viewmatrix, projmatrix):
    """markVisible (rasterize_points.cu:237-256): bool (P,) of Gaussians in front of the near plane."""
    lib = _lib.load()
    means3D = _prep(means3D, "means3D")
    P = means3D.size(0)
    present = torch.zeros((P,), dtype=torch.bool, device=means3D.device)
    if P:
        _lib.check(lib.fdls_mark_visible(P, means3D.data_ptr(), _prep(viewmatrix, "viewmatrix").data_ptr(),
                                         _prep(projmatrix, "projmatrix").data_ptr(), present.data_ptr(),
                                         _stream(means3D.device)),
                   "fdls_mark_visible")
    return present


def dist_cuda2(points):
    """simple_knn._C.distCUDA2 (call site scene/gaussian_model.py:203): (N,3) -> (N,) mean squared 3-NN distance."""
    lib = _lib.load()
    points = _prep(points, "points")
    N = points.size(0)
    out = torch.empty((N,), dtype=torch.float32, device=points.device)
    if N:
        wb = lib.fdls_knn_workspace_bytes(N)
        ws = torch.empty(wb, dtype=torch.uint8, device=points.device)
        _lib.check(lib.fdls_knn_dist2(N, points.data_ptr(), out.data_ptr(), ws.data_ptr(), wb, _stream(points.device)),
                   "fdls_knn_dist2")
    return out
