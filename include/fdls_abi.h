/* fdls_abi.h -- C ABI of the B200-native (sm_100a) differentiable Gaussian rasterizer for 4D LangSplat.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++ types, no torch types, no exceptions, no
 * device allocation inside the library.  Every entry point replaces one function of the reference's native
 * surface (paths relative to /root/reference/submodules/4d-langsplat-rasterization/):
 *
 *   fdls_forward / fdls_preprocess + fdls_bin_and_render
 *        <- CudaRasterizer::Rasterizer::forward   (cuda_rasterizer/rasterizer.h:34-61, rasterizer_impl.cu:198-348)
 *           bound to Python by RasterizeGaussiansCUDA (rasterize_points.cu:35-129, ext.cpp:16)
 *   fdls_backward
 *        <- CudaRasterizer::Rasterizer::backward  (rasterizer.h:63-95, rasterizer_impl.cu:352-469)
 *           bound by RasterizeGaussiansBackwardCUDA (rasterize_points.cu:131-235, ext.cpp:17)
 *   fdls_mark_visible
 *        <- CudaRasterizer::Rasterizer::markVisible (rasterizer.h:26-31, rasterizer_impl.cu:141-153)
 *           bound by markVisible (rasterize_points.cu:237-256, ext.cpp:18)
 *   fdls_knn_dist2
 *        <- simple_knn._C.distCUDA2 (submodule absent from the checkout; call site scene/gaussian_model.py:203)
 *
 * All device pointers are fp32 / int32 / uint8 arrays in the reference's layouts (row-major (P,3), (P,M,3), (P,F),
 * planar (C,H,W) images; 4x4 matrices in the reference's "transposed" storage, auxiliary.h:58-66).  A null pointer
 * means "absent" exactly where the reference tests data_ptr()==nullptr (forward.cu:205,241; backward.cu:425,429).
 * All work is enqueued on the cudaStream_t passed in (as void*); nothing synchronises unless documented.
 * Return value: 0 on success, negative FDLS_E* otherwise; fdls_strerror() names it.
 */
#ifndef FDLS_ABI_H
#define FDLS_ABI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FDLS_OK 0
#define FDLS_EINVAL (-1)      /* bad argument combination (mirrors the Python-side checks, __init__.py:202-206) */
#define FDLS_EWORKSPACE (-2)  /* a caller-provided workspace is too small */
#define FDLS_ECUDA (-3)       /* a CUDA launch or API call failed; see cudaGetLastError */
#define FDLS_EUNSUPPORTED (-4) /* e.g. F > FDLS_MAX_LANG_CHANNELS, or NUM_CHANNELS != 3 without colours */
#define FDLS_ELIMIT (-5)      /* size beyond the fixed maximum of an allocation-free entry point */

#define FDLS_MAX_LANG_CHANNELS 6 /* the reference is built with 3 (CLIP) or 6 (video) latents: scripts/train_eval.sh:3-4 */
#define FDLS_TILE 16             /* cuda_rasterizer/config.h:18-19 */

/* Byte sizes of the three caller-owned workspaces (the reference's geomBuffer / imgBuffer / binningBuffer,
 * rasterizer_impl.cu:155-194).  binning_bytes is for `capacity` duplicated instances. */
typedef struct FdlsWorkspaceSizes {
  size_t geom_bytes;    /* per-Gaussian state, function of P */
  size_t img_bytes;     /* per-pixel + per-tile state, function of W,H */
  size_t binning_bytes; /* keys/values double buffers + sort scratch, function of capacity */
} FdlsWorkspaceSizes;

/* Element offsets (bytes) of the arrays inside the workspaces, so tests can read intermediates bit-exactly. */
typedef struct FdlsLayout {
  /* geom */
  size_t rec;           /* float[16] per Gaussian: x,y, conic a,b,c, opacity, depth, r,g,b, lang[6] */
  size_t tiles_touched; /* u32[P] */
  size_t point_offsets; /* u32[P] inclusive scan of tiles_touched taken in DEPTH order (see depth_order) */
  size_t clamped;       /* u8[P]: bit c set if colour channel c was clamped (forward.cu:66-69) */
  size_t num_rendered;  /* u32[4]: [0]=R (device copy), [1]=overflow flag, [2..3] reserved */
  size_t scan_scratch;
  size_t depth_order;   /* u32[P]: Gaussian indices sorted (stably) by depth bits; culled ones last */
  size_t depth_scratch; /* scratch of that P-element sort */
  /* img */
  size_t final_T;   /* f32[H*W]  (reference "accum_alpha") */
  size_t n_contrib; /* u32[H*W] */
  size_t ranges;    /* uint2[tiles] */
  size_t tile_max_contrib; /* u32[tiles]: max n_contrib over the tile's pixels (lets backward skip the tail) */
  size_t tile_start;       /* u32[tiles+1]: exclusive scan of the per-tile instance counts (binning_algo 0) */
  size_t tile_cursor;      /* u32[tiles]: per-tile instance totals (binning_algo 0).  fdls_backward reuses it: afterwards it
                              holds the tiles ordered deepest-first, the launch order of the compositor backward (so two
                              fdls_backward calls on the SAME image workspace must not overlap on the GPU) */
  /* binning */
  size_t keys_alt; /* u64[capacity] ping-pong partner of `keys` */
  size_t keys;     /* u64[capacity] sorted keys after the forward (reference: point_list_keys) */
  size_t vals_alt; /* u32[capacity] */
  size_t vals;     /* u32[capacity] sorted Gaussian ids (reference: point_list) */
  size_t sort_scratch;
  size_t part_table; /* u32[ctas(capacity)][tiles]: per-CTA per-tile instance counts / prefixes (binning_algo 0) */
} FdlsLayout;

typedef struct FdlsFwdArgs {
  /* sizes */
  int32_t P;  /* Gaussians */
  int32_t D;  /* active SH degree */
  int32_t M;  /* SH coefficients per channel stored in `shs` */
  int32_t F;  /* language-feature channels (run-time; the reference bakes it in: setup.py:18,32) */
  int32_t W, H;
  /* scalars */
  float tanfovx, tanfovy, scale_modifier;
  int32_t prefiltered, debug, include_feature;
  /* inputs (device) */
  const float* background;       /* (3) */
  const float* means3D;          /* (P,3) */
  const float* shs;              /* (P,M,3) or null */
  const float* colors_precomp;   /* (P,3) or null */
  const float* language_feature; /* (P,F) or null (only when !include_feature) */
  const float* opacities;        /* (P) */
  const float* scales;           /* (P,3) or null */
  const float* rotations;        /* (P,4) or null */
  const float* cov3D_precomp;    /* (P,6) or null */
  const float* viewmatrix;       /* (16) */
  const float* projmatrix;       /* (16) */
  const float* campos;           /* (3) */
  /* outputs (device) */
  float* out_color;    /* (3,H,W) */
  float* out_language; /* (F,H,W) or null when !include_feature */
  float* out_depth;    /* (1,H,W) */
  int32_t* radii;      /* (P) */
  /* workspaces (device, 256-byte aligned) */
  void* geom_ws;    size_t geom_bytes;
  void* img_ws;     size_t img_bytes;
  void* binning_ws; size_t binning_bytes;
  int64_t capacity; /* instances the binning workspace was sized for */
  /* optional pinned, DEVICE-ACCESSIBLE host uint32[2] (cudaHostAlloc / torch pin_memory under UVA): the scan kernel
   * stores {R, overflow} into it directly; valid once the stream has passed fdls_preprocess's work */
  uint32_t* num_rendered_host;
  /* host-known instance count for grid sizing in fdls_bin_and_render; <0 = unknown, size grids by capacity.
   * <0 also states that `capacity` is the one fdls_preprocess ran with (its clipped tile ranges are reused);
   * pass the exact R (>= 0) whenever stage 2 runs with a different capacity than stage 1. */
  int64_t num_rendered_hint;
  /* optional: array of FDLS_FWD_EVENTS cudaEvent_t handles recorded on the stream between stages
   * [0] start, [1] preprocess, [2] scan, [3] duplicate, [4] sort, [5] tile ranges, [6] compositor; null = off */
  void** stage_events;
  /* binning algorithm: 0 = tile-bucketed count/scan/scatter + per-tile shared-memory sort (default),
   * 1 = key duplication + onesweep LSD radix sort (the reference's structure).  Results are identical. */
  int32_t binning_algo;
  /* binning_algo 0 never needs (tile|depth) keys: its scatter places every instance by rank.  Non-zero makes it ALSO
   * write the reference-format sorted keys (point_list_keys, rasterizer_impl.cu:307-312) into the binning workspace,
   * for parity checks; binning_algo 1 always produces them. */
  int32_t emit_keys;
} FdlsFwdArgs;
#define FDLS_FWD_EVENTS 7

typedef struct FdlsBwdArgs {
  int32_t P, D, M, F, W, H;
  float tanfovx, tanfovy, scale_modifier;
  int32_t debug, include_feature;
  int64_t num_rendered; /* R of the forward that filled the workspaces */
  const float* background;
  const float* means3D;
  const float* shs;
  const float* colors_precomp;
  const float* language_feature;
  const float* scales;
  const float* rotations;
  const float* cov3D_precomp;
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  const int32_t* radii;
  /* forward state */
  const void* geom_ws; const void* img_ws; const void* binning_ws; int64_t capacity;
  /* upstream gradients */
  const float* dL_dout_color;    /* (3,H,W) */
  const float* dL_dout_language; /* (F,H,W) or null */
  const float* dL_dout_depth;    /* (1,H,W) or null (treated as zeros) */
  /* scratch: float[16*P] per-Gaussian screen-space gradient records, zeroed by the library */
  void* grad_ws; size_t grad_bytes;
  /* gradient outputs (device); every array is fully written (zeros for invisible Gaussians); null = not wanted */
  float* dL_dmeans3D;  /* (P,3) */
  float* dL_dmeans2D;  /* (P,3), .z = 0 */
  float* dL_dsh;       /* (P,M,3) */
  float* dL_dcolors;   /* (P,3) */
  float* dL_dlanguage; /* (P,F) */
  float* dL_dopacity;  /* (P) */
  float* dL_dscales;   /* (P,3) */
  float* dL_drotations;/* (P,4) */
  float* dL_dcov3D;    /* (P,6) */
  float* dL_dconic;    /* (P,4) optional (reference returns it internally only) */
  float* dL_ddepths;   /* (P) optional */
  /* optional: FDLS_BWD_EVENTS cudaEvent_t handles: [0] start, [1] zero-fill, [2] compositor bwd, [3] preprocess bwd */
  void** stage_events;
  /* compositor backward: 0 = tensor-core reduction of the per-pixel terms (render_bwd.cu, default),
   * 1 = round-1 warp-shuffle butterfly (render_bwd_shfl.cu), kept for A/B measurements.  Same results to rounding. */
  int32_t bwd_algo;
  /* non-zero: ADD this view's gradients to the output arrays instead of overwriting them (the caller zero-fills
   * before the first view).  Lets a rank sum the views of its shard directly inside its all-reduce bucket
   *  -- no autograd accumulation pass, no pack copy.  The reference always overwrites. */
  int32_t accumulate;
  /* optional (P,3), always overwritten: dL_dRGB after the clamp mask (backward.cu:53-56).  With `shs` given, the SH
   * gradient of ONE view is the outer product of the 16 basis values of the view direction and this vector, so a
   * multi-view / multi-GPU step can keep 3 floats per view and Gaussian and expand them once with
   * fdls_sh_grad_expand instead of writing, accumulating and all-reducing 3*M floats per Gaussian; pass dL_dsh = null. */
  float* dL_dsh_rgb;
  /* non-zero (with dL_dsh_rgb): the library also writes the slot header fdls_sh_grad_expand reads -- the view's camera
   * position and the valid flag 1.0f -- into the four floats directly in FRONT of dL_dsh_rgb, so the caller does not
   * have to launch copies for it. */
  int32_t sh_rgb_header;
} FdlsBwdArgs;
#define FDLS_BWD_EVENTS 4

int fdls_abi_version(void);
/* number of kernels this library has launched in this process so far (bench.py reports the delta as gpu_launches) */
uint64_t fdls_kernel_launches(void);
const char* fdls_strerror(int code);

int fdls_workspace_sizes(int32_t P, int32_t W, int32_t H, int64_t capacity, FdlsWorkspaceSizes* out);
int fdls_workspace_layout(int32_t P, int32_t W, int32_t H, int64_t capacity, FdlsLayout* out);
size_t fdls_grad_workspace_bytes(int32_t P);

/* Stage 1: preprocess + tile-count scan.  Writes radii, the geometry workspace and R (device + optional pinned host). */
int fdls_preprocess(const FdlsFwdArgs* a, void* stream);
/* Stage 2: key duplication, radix sort, tile ranges, compositing.  Instances beyond `capacity` are dropped and the
 * overflow flag is raised (caller re-runs stage 2 with a larger binning workspace). */
int fdls_bin_and_render(const FdlsFwdArgs* a, void* stream);
/* Both stages back to back (no host synchronisation). */
int fdls_forward(const FdlsFwdArgs* a, void* stream);

int fdls_backward(const FdlsBwdArgs* a, void* stream);

/* dL_dsh (P,M,3) (+)= sum over B view slots of basis(normalize(means3D - campos_v)) (x) dL_dRGB_v (degree D).
 * `slots`: B records of `stride` floats: [0..2] camera position of the view, [3] 1.0f if the slot holds a view (0 =
 * skip), [4 .. 4+3P) the (P,3) array fdls_backward wrote through dL_dsh_rgb.  Replaces, for a batch, the per-view
 * dL_dsh writes of computeColorFromSH backward (backward.cu:41-126); the slots are what crosses NVLink. */
int fdls_sh_grad_expand(int32_t P, int32_t D, int32_t M, int32_t B, const float* means3D, const float* slots,
                        size_t stride, float* dL_dsh, int32_t accumulate, void* stream);

int fdls_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                      uint8_t* present, void* stream);

size_t fdls_knn_workspace_bytes(int32_t N);
int fdls_knn_dist2(int32_t N, const float* xyz, float* out_dist2, void* ws, size_t ws_bytes, void* stream);

/* Stable LSD radix sort of (u64 key, u32 value) pairs on key bits [begin_bit, end_bit); exported for the parity tests
 * (replaces cub::DeviceRadixSort::SortPairs, rasterizer_impl.cu:307-312).  Double-buffered like cub::DoubleBuffer:
 * input in (keys_a, vals_a); returns 0 if the sorted result is in the `a` buffers, 1 if in the `b` buffers
 * (= number of 8-bit passes modulo 2), negative on error.  n_dev may be null (then n = n_max); otherwise the
 * element count is min(*n_dev, n_max), read on the device. */
size_t fdls_sort_scratch_bytes(int64_t n_max);
int fdls_sort_pairs(uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, int64_t n_max,
                    const uint32_t* n_dev, int begin_bit, int end_bit, void* scratch, size_t scratch_bytes,
                    void* stream);
/* fdls_sort_pairs without a scratch argument and without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  The scratch is a static device array inside the library, sized for
 * n_max <= FDLS_SORT_STATIC_MAX_PAIRS (8.0 MiB, held by every GPU that loads the library until exit); larger n_max
 * returns FDLS_ELIMIT.  Same arguments, result and return value as fdls_sort_pairs otherwise; the caller still owns
 * the key / value double buffers.  Calls no cudaMalloc / cudaFree / allocating routine.  All calls on one device
 * share the array (the part a call uses is zeroed on its stream first): calls that overlap on the GPU race, so
 * serialise them on one stream or with events.  The calling thread's current device must own `stream`. */
#define FDLS_SORT_STATIC_MAX_PAIRS 2097152
int fdls_sort_pairs_static(uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, int64_t n_max,
                           const uint32_t* n_dev, int begin_bit, int end_bit, void* stream);

/* ---- SURVEY 8f N1, first step: the HexPlane feature field that feeds the deformation network ----------------------
 * Replaces HexPlaneField.forward / get_density (scene/hexplane.py:145-170): normalize_aabb (:20-21), then per
 * resolution level the PRODUCT over the six coordinate planes (xy, xz, xt, yz, yt, zt; itertools.combinations order,
 * :76-78) of a bilinear F.grid_sample(align_corners=True, padding_mode='border') (:22-47), levels concatenated
 * (:80-106).  The reference issues 6*L grid_sample launches + 6*L products + a cat and as many backward nodes; here
 * one kernel per direction.
 *
 * Plane storage: the reference parameter is (1, C, H, W) with W = reso[first coord], H = reso[second coord]
 * (init_grid_param, :60-62).  The kernels read CHANNEL-LAST copies (H, W, C): a bilinear corner is then one 4*C-byte
 * vector instead of C scattered words.  fdls_hexplane_pack converts parameter -> channel-last (once per parameter
 * update); fdls_hexplane_unpack_grad converts the channel-last gradient back into the parameter's layout. */
#define FDLS_HEX_MAX_LEVELS 4
#define FDLS_HEX_PLANES 6
typedef struct FdlsHexArgs {
  int32_t P;      /* points */
  int32_t C;      /* channels per plane (output_coordinate_dim): 4, 8, 16 or 32 */
  int32_t L;      /* resolution levels (len(multires)), <= FDLS_HEX_MAX_LEVELS */
  int32_t reso[FDLS_HEX_MAX_LEVELS][4]; /* per level: grid resolution along x, y, z, t */
  const float* pts;   /* (P,3) world positions */
  const float* time;  /* (P,1) time coordinate, used as given (already in [-1,1] grid units, hexplane.py:150-151) */
  const float* aabb;  /* (2,3) device: aabb[0], aabb[1] of normalize_aabb */
  const float* planes[FDLS_HEX_MAX_LEVELS][FDLS_HEX_PLANES]; /* channel-last (H,W,C) copies */
  float* out;         /* forward: (P, L*C) features */
  /* backward only (all optional = may be null): */
  const float* dL_dout;  /* (P, L*C) */
  float* dL_dplanes[FDLS_HEX_MAX_LEVELS][FDLS_HEX_PLANES]; /* channel-last, ACCUMULATED into (caller zero-fills) */
  float* dL_dpts;        /* (P,3) written */
  float* dL_dtime;       /* (P,1) written */
} FdlsHexArgs;
int fdls_hexplane_forward(const FdlsHexArgs* a, void* stream);
int fdls_hexplane_backward(const FdlsHexArgs* a, void* stream);
/* (C,H,W) -> (H,W,C) and back; plain tiled transposes of a C x (H*W) matrix. */
int fdls_hexplane_pack(int32_t C, int64_t HW, const float* chw, float* hwc, void* stream);
int fdls_hexplane_unpack_grad(int32_t C, int64_t HW, const float* hwc, float* chw, void* stream);
/* The same for n <= FDLS_HEX_MAX_LEVELS*FDLS_HEX_PLANES planes of C channels in ONE launch; HW, the source and the
 * destination pointers are HOST arrays of n entries (entries with HW == 0 are skipped). */
int fdls_hexplane_pack_batch(int32_t n, int32_t C, const int64_t* HW, const float* const* chw, float* const* hwc,
                             void* stream);
int fdls_hexplane_unpack_grad_batch(int32_t n, int32_t C, const int64_t* HW, const float* const* hwc,
                                    float* const* chw, void* stream);

/* ---- SURVEY 8f N1, second step: the MLP half of the deformation network on the tcgen05 tensor cores -------------
 * Y[rows x N] = act(X[rows x K] W[N x K]^T + bias), fp32 in and out, W in torch's nn.Linear layout (out, in).
 * The product runs as fp16 x 2 operand splits (three tcgen05.mma.kind::f16 terms, fp32 accumulation in TMEM), which
 * reproduces the fp32 result to ~1e-7 relative.  K % 16 == 0, N % 8 == 0, both <= 128.  The building block of
 * fdls_deform_forward; exported so that the descriptor conventions are pinned by a test of their own. */
int fdls_umma_linear(int32_t rows, int32_t N, int32_t K, const float* X, const float* W, const float* bias, float* Y,
                     int32_t relu, void* stream);

/* The deformation MLP, fused (Deformation.query_time + forward_dynamic, scene/deformation.py:70-182, with the
 * activations gaussian_renderer/__init__.py:191-193 applies afterwards):
 *     hidden = feature_out(grid_feature)                       Linear(G, W)                       (:45-58, defor_depth <= 1)
 *     head h: delta_h = Linear(W, k_h)(ReLU(Linear(W, W)(ReLU(hidden))))                           (:59-65)
 *     out_h  = act_h(base_h + delta_h)                          base = the canonical attribute      (:118-147)
 * for up to FDLS_DEFORM_MAX_HEADS heads (pos / scales / rotations / opacity / shs / discrete-coefficient heads; which
 * ones exist is the caller's choice: the reference switches them with no_dx, no_ds, no_dr, no_do, no_dshs).
 * One persistent kernel: 128 points per CTA tile, every GEMM on tcgen05 (fp16 x 2), accumulators in TMEM, activations
 * kept in shared memory between layers; the weights stream from L2 as pre-split operand chunks (fdls_deform_prepare,
 * once per parameter update).  G % 16 == 0, G <= 128, W in {64, 128}, k_h <= 48. */
#define FDLS_DEFORM_MAX_HEADS 6
#define FDLS_DEFORM_ACT_NONE 0
#define FDLS_DEFORM_ACT_EXP 1       /* scaling_activation = torch.exp (scene/gaussian_model.py:48) */
#define FDLS_DEFORM_ACT_NORMALIZE 2 /* rotation_activation = F.normalize: x / max(||x||, 1e-12) (:55) */
#define FDLS_DEFORM_ACT_SIGMOID 3   /* opacity_activation = torch.sigmoid (:52) */
typedef struct FdlsDeformHead {
  int32_t k;         /* output width of the head (3, 3, 4, 1, 48, ...) */
  int32_t act;       /* FDLS_DEFORM_ACT_* applied to base + delta */
  const float* w1;   /* (W, W)  Linear weight, torch layout (out, in) */
  const float* b1;   /* (W) */
  const float* w2;   /* (k, W) */
  const float* b2;   /* (k) */
  const float* base; /* (P, k) canonical attribute added to the head's output, or null */
  float* out;        /* (P, k) */
} FdlsDeformHead;
typedef struct FdlsDeformArgs {
  int32_t P, G, W, H;
  const float* feat; /* (P, G) grid features (fdls_hexplane_forward's output) */
  const float* w0;   /* (W, G) feature_out[0].weight */
  const float* b0;   /* (W) */
  FdlsDeformHead head[FDLS_DEFORM_MAX_HEADS];
  void* prep;        /* fdls_deform_prep_bytes() bytes, filled by fdls_deform_prepare, read by fdls_deform_forward */
  size_t prep_bytes;
} FdlsDeformArgs;
size_t fdls_deform_prep_bytes(const FdlsDeformArgs* a);
/* split every weight matrix into fp16 x 2 operand chunks (reads w0, w1, w2 of every head; P, feat, base, out unused) */
int fdls_deform_prepare(const FdlsDeformArgs* a, void* stream);
/* reads feat, biases, base and `prep`; writes out of every head */
int fdls_deform_forward(const FdlsDeformArgs* a, void* stream);

/* ---- the streaming kernels around the rasterizer in a training iteration ----------------------
 * fdls_masked_l1: *loss += weight * sum |pred*mask - gt*mask| and, if grad != null, grad = weight * sign(pred*mask -
 *   gt*mask) * mask in the same pass (l1_loss of utils/loss_utils.py:20-21 as train.py:283-296 uses it; weight = the
 *   mean's 1/N times args.lam; mask (HW floats, shared by the C channels) may be null = all ones).  pred, gt, grad:
 *   (C, HW) planar.  The gradient is exactly what fdls_backward takes as dL_dout_color / dL_dout_language.
 * fdls_gt_feature_gather: out (F, HW) = feature_map[seg[pix]] (negative ids index from the end like tensor indexing),
 *   mask[pix] = seg[pix] != -1: the per-pixel gather of Camera.get_language_feature (scene/cameras.py:97-118), which
 *   the reference does on the CPU per view per iteration.
 * fdls_adam_step: torch.optim.Adam (no weight decay / amsgrad; scene/gaussian_model.py:301) over `count` <= 32 tensors
 *   in ONE launch; p, g, m, v, n, lr, eps, step are HOST arrays of `count` entries (step = the 1-based step count of
 *   each tensor's state).
 * fdls_densify_stats: add_densification_stats (scene/gaussian_model.py:746-748).
 * fdls_compact_rows: out[dest[i]] = in[i] for rows with keep[i] (prune_points / _prune_optimizer, :487-520); dest = an
 *   exclusive scan of keep, shared by all arrays of the model. */
int fdls_masked_l1(int32_t C, int64_t HW, const float* pred, const float* gt, const float* mask, float weight, float* grad,
                   float* loss, void* stream);
int fdls_gt_feature_gather(int64_t HW, int32_t K, int32_t F, const int32_t* seg, const float* feature_map, float* out,
                           uint8_t* mask, void* stream);
int fdls_adam_step(int32_t count, float* const* p, const float* const* g, float* const* m, float* const* v,
                   const int64_t* n, const float* lr, const float* eps, const int32_t* step, double beta1, double beta2,
                   void* stream);
int fdls_densify_stats(int32_t P, const float* viewspace_grad, const uint8_t* filter, float* xyz_gradient_accum,
                       float* denom, void* stream);
int fdls_compact_rows(int64_t n, int32_t width, const uint8_t* keep, const int32_t* dest, const float* in, float* out,
                      void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FDLS_ABI_H */
