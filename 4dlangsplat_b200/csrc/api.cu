// api.cu -- the extern "C" entry points declared in include/fdls_abi.h: argument checks, workspace carving, and the
// launch sequences that replace CudaRasterizer::Rasterizer::forward / backward / markVisible
// (cuda_rasterizer/rasterizer_impl.cu:198-348, :352-469, :141-153).  No device allocation, no host synchronisation.
#include <atomic>
#include <cstdio>
#include "common.cuh"

namespace fdls {
static std::atomic<unsigned long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }
static inline void rec(void** ev, int i, cudaStream_t st) {
  if (ev && ev[i]) cudaEventRecord((cudaEvent_t)ev[i], st);
}
// preprocess_fwd.cu
cudaError_t launch_preprocess_fwd(const FdlsFwdArgs* a, const ViewParams& vp, float4* rec, uint32_t* tiles_touched,
                                  uint8_t* clamped, uint64_t* depth_keys, uint32_t* depth_vals, uint32_t* depth_hist,
                                  uint32_t* tickets, cudaStream_t st);
cudaError_t launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t st);
// binning.cu
size_t scan_scratch_bytes(int P);
size_t scan_state_bytes(int P);
cudaError_t launch_scan(int P, const uint32_t* tiles_touched, const uint32_t* order, uint32_t* offsets,
                        uint32_t* block_sums, unsigned long long* state, uint32_t* num_rendered, long long capacity,
                        uint32_t* host_mirror, cudaStream_t st);
cudaError_t launch_duplicate(int P, const float4* rec, const uint32_t* order, const uint32_t* offsets, const int* radii,
                             uint64_t* keys, uint32_t* vals, int gx, int gy, long long capacity, uint32_t* global_hist,
                             int passes, cudaStream_t st);
cudaError_t launch_tile_ranges(const uint32_t* num_rendered, long long n_grid, long long capacity, const uint64_t* keys,
                               uint2* ranges, int tiles, cudaStream_t st);
// radix_sort.cu
size_t sort_scratch_bytes(long long n_max);
int sort_pairs(uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, long long n_max,
               const uint32_t* n_dev, int begin_bit, int end_bit, void* scratch, size_t scratch_bytes,
               cudaStream_t st);
int sort_pairs_static(uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, long long n_max,
                      const uint32_t* n_dev, int begin_bit, int end_bit, cudaStream_t st);
int sort_prepare(long long n_max, int begin_bit, int end_bit, void* scratch, size_t scratch_bytes, cudaStream_t st,
                 size_t zero_before = 0);
uint32_t* sort_hist(void* scratch);
int sort_run(uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, long long n_max,
             const uint32_t* n_dev, int begin_bit, int end_bit, void* scratch, size_t scratch_bytes, bool have_hist,
             cudaStream_t st);
// binning_partition.cu
bool partition_supported(int gx, int gy);
size_t partition_table_bytes(long long capacity, int P, int T);
cudaError_t launch_partition(int P, int gx, int gy, const uint32_t* order, const uint32_t* offsets, const int* radii,
                             const float4* rec, uint32_t* table, uint32_t* tile_total, uint32_t* tile_start,
                             uint2* ranges, uint32_t* num_rendered, uint32_t* point_list, uint64_t* keys,
                             const uint32_t* depth_hist, long long capacity, cudaStream_t st);
// render_fwd.cu / render_bwd.cu / preprocess_bwd.cu
cudaError_t launch_render_fwd(int F, const uint2* ranges, const uint32_t* point_list, const float4* rec, int W, int H,
                              int gx, int gy, const float* bg, float* final_T, uint32_t* n_contrib, uint32_t* tile_max,
                              float* out_color, float* out_lang, float* out_depth, cudaStream_t st);
cudaError_t launch_render_bwd(int F, const uint2* ranges, const uint32_t* point_list, const float4* rec, int W, int H,
                              int gx, int gy, const float* bg, const float* final_T, const uint32_t* n_contrib,
                              const uint32_t* tile_max, const float* dpix, const float* dpix_F, const float* dpix_D,
                              float* grad_rec, const uint32_t* tile_order, cudaStream_t st);
cudaError_t launch_bwd_prepare(float* grad_rec, size_t n_floats, const uint32_t* tile_max, int tiles, uint32_t* tile_order,
                               cudaStream_t st);
cudaError_t launch_render_bwd_shfl(int F, const uint2* ranges, const uint32_t* point_list, const float4* rec, int W,
                                   int H, int gx, int gy, const float* bg, const float* final_T,
                                   const uint32_t* n_contrib, const uint32_t* tile_max, const float* dpix,
                                   const float* dpix_F, const float* dpix_D, float* grad_rec, cudaStream_t st);
cudaError_t launch_preprocess_bwd(const FdlsBwdArgs* a, const ViewParams& vp, const uint8_t* clamped,
                                  const float4* grad_rec, cudaStream_t st);
cudaError_t launch_sh_grad_expand(int P, int D, int M, int B, const float* means3D, const float* slots, size_t stride,
                                  float* dL_dsh, int accumulate, cudaStream_t st);
// deform_mlp.cu
int umma_linear(int rows, int N, int K, const float* X, const float* W, const float* bias, float* Y, int relu,
                cudaStream_t st);
size_t deform_prep_bytes(const FdlsDeformArgs* a);
int deform_prepare(const FdlsDeformArgs* a, cudaStream_t st);
int deform_forward(const FdlsDeformArgs* a, cudaStream_t st);
// train_ops.cu
int masked_l1(int C, long long HW, const float* pred, const float* gt, const float* mask, float weight, float* grad,
              float* loss, cudaStream_t st);
int gt_gather(long long HW, int K, int F, const int32_t* seg, const float* fmap, float* out, uint8_t* mask,
              cudaStream_t st);
int adam_step(int count, float* const* p, const float* const* g, float* const* m, float* const* v, const long long* n,
              const float* lr, const float* eps, const int* step, double b1, double b2, cudaStream_t st);
int densify_stats(int P, const float* vgrad, const uint8_t* filter, float* accum, float* denom, cudaStream_t st);
int compact_rows(long long n, int width, const uint8_t* keep, const int32_t* dest, const float* in, float* out,
                 cudaStream_t st);
// knn.cu
size_t knn_workspace_bytes(int N);
int knn_dist2(int N, const float* xyz, float* out, void* ws, size_t ws_bytes, cudaStream_t st);
int hexplane_forward(const FdlsHexArgs* a, cudaStream_t st);
int hexplane_backward(const FdlsHexArgs* a, cudaStream_t st);
int hexplane_transpose(bool pack, int n, int C, const int64_t* HW, const float* const* in, float* const* out,
                       cudaStream_t st);

// rasterizer_impl.cu:35-50
static uint32_t higher_msb(uint32_t n) {
  uint32_t msb = sizeof(n) * 4;
  uint32_t step = msb;
  while (step > 1) {
    step /= 2;
    if (n >> msb) msb += step;
    else msb -= step;
  }
  if (n >> msb) msb++;
  return msb;
}

static void compute_layout(int P, int W, int H, long long capacity, FdlsLayout* L, FdlsWorkspaceSizes* S) {
  const int gx = cdiv(W, TILE), gy = cdiv(H, TILE);
  const size_t tiles = (size_t)gx * gy, HW = (size_t)W * H, Pz = (size_t)(P < 1 ? 1 : P);
  const size_t cap = (size_t)(capacity < 1 ? 1 : capacity);
  size_t o = 0;
  L->rec = o;           o += align_up(Pz * REC_FLOATS * sizeof(float));
  L->tiles_touched = o; o += align_up(Pz * 4);
  L->point_offsets = o; o += align_up(Pz * 4);
  L->clamped = o;       o += align_up(Pz);
  L->num_rendered = o;  o += align_up(16);
  L->scan_scratch = o;  o += scan_scratch_bytes((int)Pz);
  L->depth_order = o;   o += align_up(Pz * 4);
  // [keys_a | keys_b | vals_b | scan state words | onesweep scratch]: see depth_sort_scratch()
  L->depth_scratch = o; o += align_up(Pz * 8) * 2 + align_up(Pz * 4) + scan_state_bytes((int)Pz) + sort_scratch_bytes((long long)Pz);
  S->geom_bytes = o;
  o = 0;
  L->final_T = o;          o += align_up(HW * 4);
  L->n_contrib = o;        o += align_up(HW * 4);
  L->ranges = o;           o += align_up(tiles * 8);
  L->tile_max_contrib = o; o += align_up(tiles * 4);
  L->tile_start = o;       o += align_up((tiles + 1) * 4);
  L->tile_cursor = o;      o += align_up(tiles * 4);
  S->img_bytes = o;
  o = 0;
  L->vals = o;         o += align_up(cap * 4);  // first: the backward finds point_list without knowing capacity
  L->vals_alt = o;     o += align_up(cap * 4);
  L->keys = o;         o += align_up(cap * 8);
  L->keys_alt = o;     o += align_up(cap * 8);
  L->sort_scratch = o; o += sort_scratch_bytes((long long)cap);
  L->part_table = o;   o += partition_supported(gx, gy) ? partition_table_bytes((long long)cap, (int)Pz, (int)tiles) : 256;
  S->binning_bytes = o;
}

// the depth sort's onesweep scratch inside the geometry workspace; the scan's state words sit directly in front of it
// so that ONE memset per frame (sort_prepare) zeroes both
static char* depth_sort_scratch(char* g, const FdlsLayout& L, size_t Pz) {
  return g + L.depth_scratch + 2 * align_up(Pz * 8) + align_up(Pz * 4) + scan_state_bytes((int)Pz);
}

static ViewParams make_view(const float* view, const float* proj, const float* campos, const float* bg, int W, int H,
                            float tanfovx, float tanfovy, float scale_modifier) {
  ViewParams vp;
  vp.view = view; vp.proj = proj; vp.campos = campos; vp.bg = bg;
  vp.tanfovx = tanfovx; vp.tanfovy = tanfovy;
  vp.focal_y = H / (2.0f * tanfovy);  // rasterizer_impl.cu:226-227
  vp.focal_x = W / (2.0f * tanfovx);
  vp.scale_modifier = scale_modifier;
  vp.W = W; vp.H = H; vp.gx = cdiv(W, TILE); vp.gy = cdiv(H, TILE);
  return vp;
}

static int check_fwd(const FdlsFwdArgs* a) {
  if (!a || a->P < 0 || a->W <= 0 || a->H <= 0) return FDLS_EINVAL;
  if (a->F < 0 || a->F > FDLS_MAX_LANG_CHANNELS) return FDLS_EUNSUPPORTED;
  if (a->P > 0) {
    if (!a->means3D || !a->opacities || !a->viewmatrix || !a->projmatrix || !a->radii) return FDLS_EINVAL;
    if ((a->shs == nullptr) == (a->colors_precomp == nullptr)) return FDLS_EINVAL;  // __init__.py:202-203
    if (a->shs && (a->M < (a->D + 1) * (a->D + 1) || a->D > 3 || !a->campos)) return FDLS_EINVAL;
    const bool sr = a->scales != nullptr && a->rotations != nullptr;
    if (sr == (a->cov3D_precomp != nullptr)) return FDLS_EINVAL;  // __init__.py:205-206
    if ((a->scales == nullptr) != (a->rotations == nullptr)) return FDLS_EINVAL;
  }
  if (a->include_feature && (a->F < 1 || !a->language_feature || !a->out_language)) return FDLS_EINVAL;
  if (!a->out_color || !a->out_depth || !a->background) return FDLS_EINVAL;
  if (!a->geom_ws || !a->img_ws) return FDLS_EINVAL;
  return FDLS_OK;
}

}  // namespace fdls

using namespace fdls;

extern "C" {

#define FDLS_API __attribute__((visibility("default")))

FDLS_API int fdls_abi_version(void) { return 1; }
FDLS_API uint64_t fdls_kernel_launches(void) { return g_launches.load(std::memory_order_relaxed); }

FDLS_API const char* fdls_strerror(int code) {
  switch (code) {
    case FDLS_OK: return "ok";
    case FDLS_EINVAL: return "invalid argument";
    case FDLS_EWORKSPACE: return "workspace too small";
    case FDLS_ECUDA: return "CUDA error";
    case FDLS_EUNSUPPORTED: return "unsupported configuration";
    case FDLS_ELIMIT: return "size exceeds the fixed maximum of an allocation-free entry point";
    default: return "unknown error";
  }
}

FDLS_API int fdls_workspace_sizes(int32_t P, int32_t W, int32_t H, int64_t capacity, FdlsWorkspaceSizes* out) {
  if (!out || P < 0 || W <= 0 || H <= 0 || capacity < 0) return FDLS_EINVAL;
  FdlsLayout L;
  compute_layout(P, W, H, capacity, &L, out);
  return FDLS_OK;
}

FDLS_API int fdls_workspace_layout(int32_t P, int32_t W, int32_t H, int64_t capacity, FdlsLayout* out) {
  if (!out || P < 0 || W <= 0 || H <= 0 || capacity < 0) return FDLS_EINVAL;
  FdlsWorkspaceSizes S;
  compute_layout(P, W, H, capacity, out, &S);
  return FDLS_OK;
}

FDLS_API size_t fdls_grad_workspace_bytes(int32_t P) { return align_up((size_t)(P < 1 ? 1 : P) * 16 * sizeof(float)); }

FDLS_API int fdls_preprocess(const FdlsFwdArgs* a, void* stream) {
  int rc = check_fwd(a);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  FdlsLayout L;
  FdlsWorkspaceSizes S;
  compute_layout(a->P, a->W, a->H, a->capacity < 0 ? 0 : a->capacity, &L, &S);
  if (a->geom_bytes < S.geom_bytes || a->img_bytes < S.img_bytes) return FDLS_EWORKSPACE;
  char* g = (char*)a->geom_ws;
  ViewParams vp = make_view(a->viewmatrix, a->projmatrix, a->campos, a->background, a->W, a->H, a->tanfovx, a->tanfovy,
                            a->scale_modifier);
  uint32_t* nr = (uint32_t*)(g + L.num_rendered);
  rec(a->stage_events, 0, st);
  // depth-sort scratch: [keys_a | keys_b | vals_b | onesweep scratch]; vals_a is depth_order itself when the pass
  // count is even (4 passes over 32 bits), so the sorted order lands in L.depth_order
  const size_t Pz = (size_t)(a->P < 1 ? 1 : a->P);
  uint64_t* dk_a = (uint64_t*)(g + L.depth_scratch);
  uint64_t* dk_b = (uint64_t*)(g + L.depth_scratch + align_up(Pz * 8));
  uint32_t* dv_b = (uint32_t*)(g + L.depth_scratch + 2 * align_up(Pz * 8));
  char* dscr = depth_sort_scratch(g, L, Pz);
  unsigned long long* scan_state = (unsigned long long*)(dscr - scan_state_bytes((int)Pz));
  uint32_t* order = (uint32_t*)(g + L.depth_order);
  // the depth sort's digit histograms are accumulated by the preprocess kernel itself (no separate pass over the keys)
  const size_t dscr_bytes = sort_scratch_bytes((long long)Pz);
  rc = sort_prepare(a->P, 0, 32, dscr, dscr_bytes, st, scan_state_bytes((int)Pz));
  if (rc < 0) return rc;
  if (launch_preprocess_fwd(a, vp, (float4*)(g + L.rec), (uint32_t*)(g + L.tiles_touched), (uint8_t*)(g + L.clamped),
                            dk_a, order, a->P > 0 ? sort_hist(dscr) : nullptr, nr + 2, st) != cudaSuccess)
    return FDLS_ECUDA;
  rec(a->stage_events, 1, st);
  // (1) order the GAUSSIANS by depth bits (stable, P elements) ...
  int sel = sort_run(dk_a, dk_b, order, dv_b, a->P, nullptr, 0, 32, dscr, dscr_bytes, true, st);
  if (sel < 0) return sel;
  if (sel != 0) return FDLS_ECUDA;  // 4 passes: result is back in the `a` buffers by construction
  // (2) ... and scan the tile counts in that order, so instances are emitted depth-sorted
  if (launch_scan(a->P, (uint32_t*)(g + L.tiles_touched), order, (uint32_t*)(g + L.point_offsets),
                  (uint32_t*)(g + L.scan_scratch), scan_state, nr, a->capacity, a->num_rendered_host, st) != cudaSuccess)
    return FDLS_ECUDA;
  rec(a->stage_events, 2, st);
  return FDLS_OK;
}

FDLS_API int fdls_bin_and_render(const FdlsFwdArgs* a, void* stream) {
  int rc = check_fwd(a);
  if (rc) return rc;
  if (a->capacity < 0 || (a->capacity > 0 && !a->binning_ws)) return FDLS_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  FdlsLayout L;
  FdlsWorkspaceSizes S;
  compute_layout(a->P, a->W, a->H, a->capacity, &L, &S);
  if (a->geom_bytes < S.geom_bytes || a->img_bytes < S.img_bytes) return FDLS_EWORKSPACE;
  if (a->capacity > 0 && a->binning_bytes < S.binning_bytes) return FDLS_EWORKSPACE;
  char* g = (char*)a->geom_ws;
  char* im = (char*)a->img_ws;
  char* b = (char*)a->binning_ws;
  const int gx = cdiv(a->W, TILE), gy = cdiv(a->H, TILE);
  uint32_t* nr = (uint32_t*)(g + L.num_rendered);
  long long n_grid = a->capacity;
  if (a->num_rendered_hint >= 0 && a->num_rendered_hint < n_grid) n_grid = a->num_rendered_hint;
  uint2* ranges = (uint2*)(im + L.ranges);
  const uint32_t* point_list = nullptr;
  if (a->binning_algo == 0 && partition_supported(gx, gy)) {
    if (n_grid > 0) {
      // digit histograms of the depth sort (top-byte bin 0xff = culled Gaussians), still in the geometry workspace
      const size_t Pz2 = (size_t)(a->P < 1 ? 1 : a->P);
      const uint32_t* depth_hist =
          sort_hist(depth_sort_scratch(g, L, Pz2));
      // binning_partition.cu: count / scans / scatter -- every instance goes straight to its final sorted slot
      if (launch_partition(a->P, gx, gy, (const uint32_t*)(g + L.depth_order), (const uint32_t*)(g + L.point_offsets),
                           a->radii, (const float4*)(g + L.rec), (uint32_t*)(b + L.part_table),
                           (uint32_t*)(im + L.tile_cursor), (uint32_t*)(im + L.tile_start), ranges, nr,
                           (uint32_t*)(b + L.vals), a->emit_keys ? (uint64_t*)(b + L.keys) : nullptr, depth_hist,
                           a->capacity,
                           st) != cudaSuccess)
        return FDLS_ECUDA;
      point_list = (const uint32_t*)(b + L.vals);
    } else {
      if (cudaMemsetAsync(ranges, 0, (size_t)gx * gy * sizeof(uint2), st) != cudaSuccess) return FDLS_ECUDA;
    }
    rec(a->stage_events, 3, st);
    rec(a->stage_events, 4, st);
  } else if (n_grid > 0) {
    // Instances were emitted in depth order, so a STABLE sort on the tile bits alone yields exactly the order of the
    // reference's sort on bits [0, 32+bit) (rasterizer_impl.cu:304-312): 2 passes over R instead of 6.
    const int end_bit = 32 + (int)higher_msb((uint32_t)(gx * gy));
    const int passes = (end_bit - 32 + 7) / 8;
    // choose the input buffer so that the sorted result lands in (keys, vals)
    uint64_t* k_in = (uint64_t*)(b + ((passes & 1) ? L.keys_alt : L.keys));
    uint64_t* k_out = (uint64_t*)(b + ((passes & 1) ? L.keys : L.keys_alt));
    uint32_t* v_in = (uint32_t*)(b + ((passes & 1) ? L.vals_alt : L.vals));
    uint32_t* v_out = (uint32_t*)(b + ((passes & 1) ? L.vals : L.vals_alt));
    void* sscr = b + L.sort_scratch;
    const size_t sscr_bytes = S.binning_bytes - L.sort_scratch;
    rc = sort_prepare(n_grid, 32, end_bit, sscr, sscr_bytes, st);
    if (rc < 0) return rc;
    if (launch_duplicate(a->P, (const float4*)(g + L.rec), (const uint32_t*)(g + L.depth_order),
                         (const uint32_t*)(g + L.point_offsets), a->radii, k_in, v_in, gx, gy,
                         n_grid < a->capacity ? n_grid : a->capacity, sort_hist(sscr), passes, st) != cudaSuccess)
      return FDLS_ECUDA;
    rec(a->stage_events, 3, st);
    rc = sort_run(k_in, k_out, v_in, v_out, n_grid, nr, 32, end_bit, sscr, sscr_bytes, true, st);
    if (rc < 0) return rc;
    rec(a->stage_events, 4, st);
    point_list = (const uint32_t*)(b + L.vals);
    if (launch_tile_ranges(nr, n_grid, a->capacity, (const uint64_t*)(b + L.keys), ranges, gx * gy, st) != cudaSuccess)
      return FDLS_ECUDA;
  } else {
    if (cudaMemsetAsync(ranges, 0, (size_t)gx * gy * sizeof(uint2), st) != cudaSuccess) return FDLS_ECUDA;
  }
  rec(a->stage_events, 5, st);
  const int F = a->include_feature ? a->F : 0;
  if (launch_render_fwd(F, ranges, point_list, (const float4*)(g + L.rec), a->W, a->H, gx, gy, a->background,
                        (float*)(im + L.final_T), (uint32_t*)(im + L.n_contrib), (uint32_t*)(im + L.tile_max_contrib),
                        a->out_color, a->out_language, a->out_depth, st) != cudaSuccess)
    return FDLS_ECUDA;
  rec(a->stage_events, 6, st);
  return FDLS_OK;
}

FDLS_API int fdls_forward(const FdlsFwdArgs* a, void* stream) {
  int rc = fdls_preprocess(a, stream);
  if (rc) return rc;
  return fdls_bin_and_render(a, stream);
}

FDLS_API int fdls_backward(const FdlsBwdArgs* a, void* stream) {
  if (!a || a->P < 0 || a->W <= 0 || a->H <= 0) return FDLS_EINVAL;
  if (a->F < 0 || a->F > FDLS_MAX_LANG_CHANNELS) return FDLS_EUNSUPPORTED;
  if (a->P == 0) return FDLS_OK;
  if (!a->geom_ws || !a->img_ws || !a->grad_ws || !a->dL_dout_color || !a->radii || !a->means3D) return FDLS_EINVAL;
  if (!a->viewmatrix || !a->projmatrix || !a->background) return FDLS_EINVAL;
  if (a->shs && !a->campos) return FDLS_EINVAL;  // like the reference, campos is only read by the SH colour path
  if (a->include_feature && (a->F < 1 || !a->dL_dout_language)) return FDLS_EINVAL;
  if (a->grad_bytes < fdls_grad_workspace_bytes(a->P)) return FDLS_EWORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  FdlsLayout L;
  FdlsWorkspaceSizes S;
  compute_layout(a->P, a->W, a->H, a->capacity < 0 ? 0 : a->capacity, &L, &S);
  const char* g = (const char*)a->geom_ws;
  const char* im = (const char*)a->img_ws;
  const char* b = (const char*)a->binning_ws;
  const int gx = cdiv(a->W, TILE), gy = cdiv(a->H, TILE);
  ViewParams vp = make_view(a->viewmatrix, a->projmatrix, a->campos, a->background, a->W, a->H, a->tanfovx, a->tanfovy,
                            a->scale_modifier);
  rec(a->stage_events, 0, st);
  // zero the gradient records; the same kernel orders the tiles deepest-first for render_bwd (the order goes into
  // tile_cursor, a scratch array of the forward's binning that is dead once the forward has run)
  static const bool use_order = [] {
    const char* e = getenv("FDLS_TILE_ORDER");
    return !(e && e[0] == '0');
  }();
  const bool render = a->num_rendered > 0 && b;
  uint32_t* tile_order = render && use_order && a->bwd_algo != 1 ? (uint32_t*)(const_cast<char*>(im) + L.tile_cursor) : nullptr;
  if (launch_bwd_prepare((float*)a->grad_ws, (size_t)a->P * 16, (const uint32_t*)(im + L.tile_max_contrib), gx * gy,
                         tile_order, st) != cudaSuccess)
    return FDLS_ECUDA;
  rec(a->stage_events, 1, st);
  if (render) {
    const int F = a->include_feature ? a->F : 0;
    cudaError_t e;
    if (a->bwd_algo == 1)
      e = launch_render_bwd_shfl(F, (const uint2*)(im + L.ranges), (const uint32_t*)(b + L.vals),
                                 (const float4*)(g + L.rec), a->W, a->H, gx, gy, a->background,
                                 (const float*)(im + L.final_T), (const uint32_t*)(im + L.n_contrib),
                                 (const uint32_t*)(im + L.tile_max_contrib), a->dL_dout_color, a->dL_dout_language,
                                 a->dL_dout_depth, (float*)a->grad_ws, st);
    else
      e = launch_render_bwd(F, (const uint2*)(im + L.ranges), (const uint32_t*)(b + L.vals), (const float4*)(g + L.rec),
                            a->W, a->H, gx, gy, a->background, (const float*)(im + L.final_T),
                            (const uint32_t*)(im + L.n_contrib), (const uint32_t*)(im + L.tile_max_contrib),
                            a->dL_dout_color, a->dL_dout_language, a->dL_dout_depth, (float*)a->grad_ws, tile_order, st);
    if (e != cudaSuccess) return FDLS_ECUDA;
  }
  rec(a->stage_events, 2, st);
  if (launch_preprocess_bwd(a, vp, (const uint8_t*)(g + L.clamped), (const float4*)a->grad_ws, st) != cudaSuccess)
    return FDLS_ECUDA;
  rec(a->stage_events, 3, st);
  return FDLS_OK;
}

FDLS_API int fdls_sh_grad_expand(int32_t P, int32_t D, int32_t M, int32_t B, const float* means3D, const float* slots,
                                 size_t stride, float* dL_dsh, int32_t accumulate, void* stream) {
  if (P < 0 || B < 0 || D < 0 || D > 3 || M < (D + 1) * (D + 1)) return FDLS_EINVAL;
  if (P > 0 && B > 0 && (!means3D || !slots || !dL_dsh || stride < (size_t)4 + 3 * (size_t)P)) return FDLS_EINVAL;
  return launch_sh_grad_expand(P, D, M, B, means3D, slots, stride, dL_dsh, accumulate, (cudaStream_t)stream) ==
                 cudaSuccess ? FDLS_OK : FDLS_ECUDA;
}

FDLS_API int fdls_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                               uint8_t* present, void* stream) {
  (void)projmatrix;  // the reference's x/y frustum test is commented out (auxiliary.h:154)
  if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) return FDLS_EINVAL;
  return launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream) == cudaSuccess ? FDLS_OK : FDLS_ECUDA;
}

FDLS_API size_t fdls_knn_workspace_bytes(int32_t N) { return knn_workspace_bytes(N); }

FDLS_API int fdls_knn_dist2(int32_t N, const float* xyz, float* out_dist2, void* ws, size_t ws_bytes, void* stream) {
  if (N < 0 || (N > 0 && (!xyz || !out_dist2 || !ws))) return FDLS_EINVAL;
  return knn_dist2(N, xyz, out_dist2, ws, ws_bytes, (cudaStream_t)stream);
}

FDLS_API size_t fdls_sort_scratch_bytes(int64_t n_max) { return sort_scratch_bytes(n_max); }

FDLS_API int fdls_sort_pairs(uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, int64_t n_max,
                             const uint32_t* n_dev, int begin_bit, int end_bit, void* scratch, size_t scratch_bytes,
                             void* stream) {
  if (n_max < 0 || (n_max > 0 && (!keys_a || !keys_b || !vals_a || !vals_b || !scratch))) return FDLS_EINVAL;
  return sort_pairs(keys_a, keys_b, vals_a, vals_b, n_max, n_dev, begin_bit, end_bit, scratch, scratch_bytes,
                    (cudaStream_t)stream);
}

FDLS_API int fdls_sort_pairs_static(uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b,
                                    int64_t n_max, const uint32_t* n_dev, int begin_bit, int end_bit, void* stream) {
  if (n_max < 0 || (n_max > 0 && (!keys_a || !keys_b || !vals_a || !vals_b))) return FDLS_EINVAL;
  return sort_pairs_static(keys_a, keys_b, vals_a, vals_b, n_max, n_dev, begin_bit, end_bit, (cudaStream_t)stream);
}

FDLS_API int fdls_hexplane_forward(const FdlsHexArgs* a, void* stream) {
  return hexplane_forward(a, (cudaStream_t)stream);
}
FDLS_API int fdls_hexplane_backward(const FdlsHexArgs* a, void* stream) {
  return hexplane_backward(a, (cudaStream_t)stream);
}
FDLS_API int fdls_hexplane_pack(int32_t C, int64_t HW, const float* chw, float* hwc, void* stream) {
  return hexplane_transpose(true, 1, C, &HW, &chw, &hwc, (cudaStream_t)stream);
}
FDLS_API int fdls_hexplane_unpack_grad(int32_t C, int64_t HW, const float* hwc, float* chw, void* stream) {
  return hexplane_transpose(false, 1, C, &HW, &hwc, &chw, (cudaStream_t)stream);
}
FDLS_API int fdls_hexplane_pack_batch(int32_t n, int32_t C, const int64_t* HW, const float* const* chw,
                                      float* const* hwc, void* stream) {
  return hexplane_transpose(true, n, C, HW, chw, hwc, (cudaStream_t)stream);
}
FDLS_API int fdls_hexplane_unpack_grad_batch(int32_t n, int32_t C, const int64_t* HW, const float* const* hwc,
                                             float* const* chw, void* stream) {
  return hexplane_transpose(false, n, C, HW, hwc, chw, (cudaStream_t)stream);
}

FDLS_API int fdls_umma_linear(int32_t rows, int32_t N, int32_t K, const float* X, const float* W, const float* bias,
                              float* Y, int32_t relu, void* stream) {
  if (rows > 0 && (!X || !W || !Y)) return FDLS_EINVAL;
  return umma_linear(rows, N, K, X, W, bias, Y, relu, (cudaStream_t)stream);
}

FDLS_API size_t fdls_deform_prep_bytes(const FdlsDeformArgs* a) { return deform_prep_bytes(a); }
FDLS_API int fdls_deform_prepare(const FdlsDeformArgs* a, void* stream) { return deform_prepare(a, (cudaStream_t)stream); }
FDLS_API int fdls_deform_forward(const FdlsDeformArgs* a, void* stream) { return deform_forward(a, (cudaStream_t)stream); }

FDLS_API int fdls_masked_l1(int32_t C, int64_t HW, const float* pred, const float* gt, const float* mask, float weight,
                            float* grad, float* loss, void* stream) {
  return masked_l1(C, HW, pred, gt, mask, weight, grad, loss, (cudaStream_t)stream);
}
FDLS_API int fdls_gt_feature_gather(int64_t HW, int32_t K, int32_t F, const int32_t* seg, const float* feature_map,
                                    float* out, uint8_t* mask, void* stream) {
  return gt_gather(HW, K, F, seg, feature_map, out, mask, (cudaStream_t)stream);
}
FDLS_API int fdls_adam_step(int32_t count, float* const* p, const float* const* g, float* const* m, float* const* v,
                            const int64_t* n, const float* lr, const float* eps, const int32_t* step, double beta1,
                            double beta2, void* stream) {
  if (count > 0 && (!p || !g || !m || !v || !n || !lr || !eps || !step)) return FDLS_EINVAL;
  return adam_step(count, p, g, m, v, (const long long*)n, lr, eps, step, beta1, beta2, (cudaStream_t)stream);
}
FDLS_API int fdls_densify_stats(int32_t P, const float* viewspace_grad, const uint8_t* filter, float* xyz_gradient_accum,
                                float* denom, void* stream) {
  return densify_stats(P, viewspace_grad, filter, xyz_gradient_accum, denom, (cudaStream_t)stream);
}
FDLS_API int fdls_compact_rows(int64_t n, int32_t width, const uint8_t* keep, const int32_t* dest, const float* in,
                               float* out, void* stream) {
  return compact_rows(n, width, keep, dest, in, out, (cudaStream_t)stream);
}

}  // extern "C"
