"""ctypes view of include/fdls_abi.h (the C-ABI shared library 4dlangsplat_b200/libfdls.so).

This is the only place that knows the struct layouts.  There is NO fallback: if the library is missing or fails to
load, importing this module raises, and every op above it fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FDLS_LIB", os.path.join(_HERE, "libfdls.so"))  # FDLS_LIB: an A/B build of the same ABI

FDLS_MAX_LANG_CHANNELS = 6
FDLS_TILE = 16
FDLS_FWD_EVENTS = 7
FDLS_BWD_EVENTS = 4

_f32p = C.c_void_p  # device pointers travel as plain addresses


class FdlsWorkspaceSizes(C.Structure):
    _fields_ = [("geom_bytes", C.c_size_t), ("img_bytes", C.c_size_t), ("binning_bytes", C.c_size_t)]


class FdlsLayout(C.Structure):
    _fields_ = [(n, C.c_size_t) for n in (
        "rec", "tiles_touched", "point_offsets", "clamped", "num_rendered", "scan_scratch",
        "depth_order", "depth_scratch",
        "final_T", "n_contrib", "ranges", "tile_max_contrib", "tile_start", "tile_cursor",
        "keys_alt", "keys", "vals_alt", "vals", "sort_scratch", "part_table")]


class FdlsFwdArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("F", C.c_int32), ("W", C.c_int32), ("H", C.c_int32),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("scale_modifier", C.c_float),
        ("prefiltered", C.c_int32), ("debug", C.c_int32), ("include_feature", C.c_int32),
        ("background", _f32p), ("means3D", _f32p), ("shs", _f32p), ("colors_precomp", _f32p),
        ("language_feature", _f32p), ("opacities", _f32p), ("scales", _f32p), ("rotations", _f32p),
        ("cov3D_precomp", _f32p), ("viewmatrix", _f32p), ("projmatrix", _f32p), ("campos", _f32p),
        ("out_color", _f32p), ("out_language", _f32p), ("out_depth", _f32p), ("radii", C.c_void_p),
        ("geom_ws", C.c_void_p), ("geom_bytes", C.c_size_t),
        ("img_ws", C.c_void_p), ("img_bytes", C.c_size_t),
        ("binning_ws", C.c_void_p), ("binning_bytes", C.c_size_t),
        ("capacity", C.c_int64),
        ("num_rendered_host", C.c_void_p),
        ("num_rendered_hint", C.c_int64),
        ("stage_events", C.POINTER(C.c_void_p)),
        ("binning_algo", C.c_int32),
        ("emit_keys", C.c_int32),
    ]


class FdlsBwdArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("F", C.c_int32), ("W", C.c_int32), ("H", C.c_int32),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("scale_modifier", C.c_float),
        ("debug", C.c_int32), ("include_feature", C.c_int32),
        ("num_rendered", C.c_int64),
        ("background", _f32p), ("means3D", _f32p), ("shs", _f32p), ("colors_precomp", _f32p),
        ("language_feature", _f32p), ("scales", _f32p), ("rotations", _f32p), ("cov3D_precomp", _f32p),
        ("viewmatrix", _f32p), ("projmatrix", _f32p), ("campos", _f32p), ("radii", C.c_void_p),
        ("geom_ws", C.c_void_p), ("img_ws", C.c_void_p), ("binning_ws", C.c_void_p), ("capacity", C.c_int64),
        ("dL_dout_color", _f32p), ("dL_dout_language", _f32p), ("dL_dout_depth", _f32p),
        ("grad_ws", C.c_void_p), ("grad_bytes", C.c_size_t),
        ("dL_dmeans3D", _f32p), ("dL_dmeans2D", _f32p), ("dL_dsh", _f32p), ("dL_dcolors", _f32p),
        ("dL_dlanguage", _f32p), ("dL_dopacity", _f32p), ("dL_dscales", _f32p), ("dL_drotations", _f32p),
        ("dL_dcov3D", _f32p), ("dL_dconic", _f32p), ("dL_ddepths", _f32p),
        ("stage_events", C.POINTER(C.c_void_p)),
        ("bwd_algo", C.c_int32),
        ("accumulate", C.c_int32),
        ("dL_dsh_rgb", _f32p),
        ("sh_rgb_header", C.c_int32),
    ]


# every symbol include/fdls_abi.h declares: name -> (restype, argtypes)
FDLS_HEX_MAX_LEVELS = 4
FDLS_HEX_PLANES = 6


class FdlsHexArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("C", C.c_int32), ("L", C.c_int32),
        ("reso", (C.c_int32 * 4) * FDLS_HEX_MAX_LEVELS),
        ("pts", _f32p), ("time", _f32p), ("aabb", _f32p),
        ("planes", (C.c_void_p * FDLS_HEX_PLANES) * FDLS_HEX_MAX_LEVELS),
        ("out", _f32p),
        ("dL_dout", _f32p),
        ("dL_dplanes", (C.c_void_p * FDLS_HEX_PLANES) * FDLS_HEX_MAX_LEVELS),
        ("dL_dpts", _f32p), ("dL_dtime", _f32p),
    ]


FDLS_DEFORM_MAX_HEADS = 6
FDLS_DEFORM_ACT_NONE, FDLS_DEFORM_ACT_EXP, FDLS_DEFORM_ACT_NORMALIZE, FDLS_DEFORM_ACT_SIGMOID = 0, 1, 2, 3


class FdlsDeformHead(C.Structure):
    _fields_ = [("k", C.c_int32), ("act", C.c_int32), ("w1", _f32p), ("b1", _f32p), ("w2", _f32p), ("b2", _f32p),
                ("base", _f32p), ("out", _f32p)]


class FdlsDeformArgs(C.Structure):
    _fields_ = [("P", C.c_int32), ("G", C.c_int32), ("W", C.c_int32), ("H", C.c_int32),
                ("feat", _f32p), ("w0", _f32p), ("b0", _f32p),
                ("head", FdlsDeformHead * FDLS_DEFORM_MAX_HEADS),
                ("prep", C.c_void_p), ("prep_bytes", C.c_size_t)]


SYMBOLS = {
    "fdls_abi_version": (C.c_int, []),
    "fdls_kernel_launches": (C.c_uint64, []),
    "fdls_strerror": (C.c_char_p, [C.c_int]),
    "fdls_workspace_sizes": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.POINTER(FdlsWorkspaceSizes)]),
    "fdls_workspace_layout": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.POINTER(FdlsLayout)]),
    "fdls_grad_workspace_bytes": (C.c_size_t, [C.c_int32]),
    "fdls_preprocess": (C.c_int, [C.POINTER(FdlsFwdArgs), C.c_void_p]),
    "fdls_bin_and_render": (C.c_int, [C.POINTER(FdlsFwdArgs), C.c_void_p]),
    "fdls_forward": (C.c_int, [C.POINTER(FdlsFwdArgs), C.c_void_p]),
    "fdls_backward": (C.c_int, [C.POINTER(FdlsBwdArgs), C.c_void_p]),
    "fdls_sh_grad_expand": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_size_t,
                                      C.c_void_p, C.c_int32, C.c_void_p]),
    "fdls_umma_linear": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_int32, C.c_void_p]),
    "fdls_deform_prep_bytes": (C.c_size_t, [C.POINTER(FdlsDeformArgs)]),
    "fdls_deform_prepare": (C.c_int, [C.POINTER(FdlsDeformArgs), C.c_void_p]),
    "fdls_deform_forward": (C.c_int, [C.POINTER(FdlsDeformArgs), C.c_void_p]),
    "fdls_masked_l1": (C.c_int, [C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p,
                                 C.c_void_p, C.c_void_p]),
    "fdls_gt_feature_gather": (C.c_int, [C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p]),
    "fdls_adam_step": (C.c_int, [C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                 C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.POINTER(C.c_float), C.POINTER(C.c_float),
                                 C.POINTER(C.c_int32), C.c_double, C.c_double, C.c_void_p]),
    "fdls_densify_stats": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fdls_compact_rows": (C.c_int, [C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fdls_mark_visible": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fdls_knn_workspace_bytes": (C.c_size_t, [C.c_int32]),
    "fdls_knn_dist2": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "fdls_sort_scratch_bytes": (C.c_size_t, [C.c_int64]),
    "fdls_sort_pairs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int,
                                  C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]),
    "fdls_sort_pairs_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int,
                                         C.c_int, C.c_void_p]),
    "fdls_hexplane_forward": (C.c_int, [C.POINTER(FdlsHexArgs), C.c_void_p]),
    "fdls_hexplane_backward": (C.c_int, [C.POINTER(FdlsHexArgs), C.c_void_p]),
    "fdls_hexplane_pack": (C.c_int, [C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fdls_hexplane_unpack_grad": (C.c_int, [C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fdls_hexplane_pack_batch": (C.c_int, [C.c_int32, C.c_int32, C.POINTER(C.c_int64), C.POINTER(C.c_void_p),
                                           C.POINTER(C.c_void_p), C.c_void_p]),
    "fdls_hexplane_unpack_grad_batch": (C.c_int, [C.c_int32, C.c_int32, C.POINTER(C.c_int64), C.POINTER(C.c_void_p),
                                                  C.POINTER(C.c_void_p), C.c_void_p]),
}

_lib = None


def load():
    """dlopen libfdls.so and bind every declared symbol.  Raises if anything is missing (no fallback path exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s not built: run `python 4dlangsplat_b200/build.py` (or __graft_entry__.build())"
                              % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(code, what):
    if code < 0:
        raise RuntimeError("%s failed: %s (%d)" % (what, load().fdls_strerror(code).decode(), code))
    return code


def workspace_sizes(P, W, H, capacity):
    s = FdlsWorkspaceSizes()
    check(load().fdls_workspace_sizes(P, W, H, capacity, C.byref(s)), "fdls_workspace_sizes")
    return s


def workspace_layout(P, W, H, capacity):
    s = FdlsLayout()
    check(load().fdls_workspace_layout(P, W, H, capacity, C.byref(s)), "fdls_workspace_layout")
    return s
