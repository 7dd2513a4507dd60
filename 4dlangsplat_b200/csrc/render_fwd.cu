// render_fwd.cu -- front-to-back alpha compositing of RGB + F language-feature channels + depth, one 16x16 tile per
// CTA, one pixel per thread.  Replaces renderCUDA<3,F> forward (cuda_rasterizer/forward.cu:261-395).
//
// Per-pixel arithmetic (power, alpha, transmittance test, accumulation order) is the reference's, expression for
// expression, so n_contrib / final_T are bit-identical and the images agree to rounding.
//
// B200 design:
//  * every Gaussian is ONE 64-byte record {x,y,conic,opacity,depth,rgb,lang[<=6]} written by preprocess; a batch of
//    256 records is gathered with 16-byte cp.async.cg (LDGSTS, L2 -> shared, no register staging) into a
//    double-buffered 2 x 16 KB shared-memory ring, so the gather of batch i+1 overlaps the blend of batch i.
//    The reference instead gathers colour / language / depth from GLOBAL memory per contributing (pixel, Gaussian)
//    pair (forward.cu:362-369).
//  * each warp owns an 8x4-pixel block; the staging thread computes an exact per-block visibility mask per Gaussian
//    (common.cuh block_summary: minimum of the conic form over the block's rectangle) and every warp compacts the
//    batch to the Gaussians that can touch its block before blending (most of a tile's list misses any given block,
//    and skipping them one by one was the dominant instruction cost).
//  * a conservative log-domain pre-test skips the expf for pairs whose alpha is certainly below 1/255; the exact
//    reference test still decides every pair near the threshold, so results do not change.
//  * F is a template parameter dispatched at run time (the reference needs a rebuild per F: setup.py:18,32).
#include "common.cuh"

namespace fdls {

constexpr int RF_THREADS = TILE_PIX;  // 256

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

#ifndef RF_MIN_CTAS
#define RF_MIN_CTAS 5
#endif
// 5 CTAs per SM: 48 registers (no spills) and 5 x 38.9 KB of shared memory
template <int F>
__global__ void __launch_bounds__(RF_THREADS, RF_MIN_CTAS)
render_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                  const float4* __restrict__ rec, int W, int H, int gx, const float* __restrict__ bg_color,
                  float* __restrict__ final_T, uint32_t* __restrict__ n_contrib, uint32_t* __restrict__ tile_max_contrib,
                  float* __restrict__ out_color, float* __restrict__ out_lang, float* __restrict__ out_depth) {
  pdl_grid_sync();
  // [buf][slot 0..3][thread]: conflict-free cp.async writes, broadcast reads
  __shared__ float4 s_rec[2][4][RF_THREADS];
  __shared__ float2 s_aux[2][RF_THREADS];  // {alpha-threshold in the log domain, block_summary()} of each staged Gaussian
  __shared__ uint8_t s_list[RF_THREADS / 32][RF_THREADS];  // per-warp compacted indices into the staged batch
  __shared__ uint32_t s_max;

  const int tid = threadIdx.x;
  const uint32_t tile = blockIdx.x;
  const uint32_t tx = tile % gx, ty = tile / gx;
  const int lane = tid & 31, warp = tid >> 5;
  // warp w owns the 8x4-pixel block bx = w & 1, by = w >> 1 of the tile (a compact footprint: fewer Gaussians reach it
  // than reach a 16x2 strip, and block_summary() tests the block exactly)
  const uint2 pix = {tx * TILE + (warp & 1) * 8 + (lane & 7), ty * TILE + (warp >> 1) * 4 + (lane >> 3)};
  const uint32_t pix_id = W * pix.y + pix.x;
  const float2 pixf = {(float)pix.x, (float)pix.y};
  const bool inside = pix.x < (uint32_t)W && pix.y < (uint32_t)H;
  bool done = !inside;
  const uint32_t strip_bit = 1u << warp;
  const uint32_t lt_mask = (1u << lane) - 1u;

  const uint2 range = ranges[tile];
  int toDo = range.y - range.x;
  const int rounds = (toDo + RF_THREADS - 1) / RF_THREADS;
  if (tid == 0) s_max = 0;
  __syncthreads();

  auto stage = [&](int round, int buf) {
    const uint32_t progress = round * RF_THREADS + tid;
    if (range.x + progress < range.y) {
      const uint32_t id = point_list[range.x + progress];
      const float4* src = rec + (size_t)id * 4;
#pragma unroll
      for (int k = 0; k < 4; ++k) cp_async16(&s_rec[buf][k][tid], src + k);
    }
    cp_async_commit();
  };

  float T = 1.0f;
  uint32_t last_contributor = 0;
  float C[3] = {0, 0, 0};
  float D = 0;
  float Fa[F > 0 ? F : 1] = {0};

  if (rounds > 0) stage(0, 0);
  for (int i = 0; i < rounds; i++, toDo -= RF_THREADS) {
    const int buf = i & 1;
    if (i + 1 < rounds) {
      stage(i + 1, buf ^ 1);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    {  // my own staged record (own cp.async data is visible to the issuing thread after the wait)
      const float4 r0 = s_rec[buf][0][tid], r1 = s_rec[buf][1][tid];
      const float x0 = (float)(tx * TILE), y0 = (float)(ty * TILE);
      const uint32_t bits = block_summary(r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, x0, y0, 0) |
                            block_summary(r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, x0, y0, 1);
      s_aux[buf][tid] = make_float2(-__logf(255.0f * r1.y) - 0.01f, __uint_as_float(bits));  // .x: see strip_summary
    }
    __syncthreads();

    const int cnt = min(RF_THREADS, toDo);
    // Per-warp compaction: the indices of the staged Gaussians that can reach this warp's 8x4 block are gathered
    // (in order) into a private list, so the blend loop never iterates over a Gaussian the whole warp would skip.
    // `contributor` is the 1-based position in the tile's range, exactly the reference's running counter.
    int n_hit = 0;
    if (!__all_sync(0xffffffffu, done)) {
#pragma unroll
      for (int k = 0; k < RF_THREADS / 32; ++k) {
        const int j = k * 32 + lane;
        const bool hit = j < cnt && (__float_as_uint(s_aux[buf][j].y) & strip_bit);
        const uint32_t m = __ballot_sync(0xffffffffu, hit);
        if (hit) s_list[warp][n_hit + __popc(m & lt_mask)] = (uint8_t)j;
        n_hit += __popc(m);
      }
      __syncwarp();
    }
    const uint32_t base = (uint32_t)i * RF_THREADS;
    for (int q = 0; !done && q < n_hit; q++) {
      const int j = s_list[warp][q];
      const float thr = s_aux[buf][j].x;
      const float4 a = s_rec[buf][0][j];  // x, y, conic.x, conic.y
      const float4 b = s_rec[buf][1][j];  // conic.z, opacity, depth, r
      const float2 d = {a.x - pixf.x, a.y - pixf.y};
      const float power = -0.5f * (a.z * d.x * d.x + b.x * d.y * d.y) - a.w * d.x * d.y;
      if (power > 0.0f) continue;
      if (power < thr) continue;  // certainly alpha < 1/255 (pre-test; exact test below)
      const float alpha = min(0.99f, b.y * exp(power));
      if (alpha < 1.0f / 255.0f) continue;
      const float test_T = T * (1 - alpha);
      if (test_T < 0.0001f) {
        done = true;
        break;
      }
      const float4 c = s_rec[buf][2][j];  // g, b, l0, l1
      // the reference forms (c * alpha) * T per channel (forward.cu:362-369); c * (alpha * T) is one multiply per pair
      // instead of one per channel and differs by one rounding (2^-24 relative per term; images are held to 1e-5)
      const float w = alpha * T;
      C[0] = fmaf(b.w, w, C[0]);
      C[1] = fmaf(c.x, w, C[1]);
      C[2] = fmaf(c.y, w, C[2]);
      if (F > 0) {
        const float4 e = s_rec[buf][3][j];  // l2..l5
        const float l[6] = {c.z, c.w, e.x, e.y, e.z, e.w};
#pragma unroll
        for (int ch = 0; ch < F; ch++) Fa[ch] = fmaf(l[ch], w, Fa[ch]);
      }
      D = fmaf(b.z, w, D);
      T = test_T;
      last_contributor = base + j + 1;
    }
    // also fences the shared ring before the next prefetch overwrites buffer `buf`
    if (__syncthreads_count(done) == RF_THREADS) break;
  }
  cp_async_wait<0>();

  const uint32_t wmax = __reduce_max_sync(0xffffffffu, inside ? last_contributor : 0u);
  if ((tid & 31) == 0) atomicMax(&s_max, wmax);
  if (inside) {
    final_T[pix_id] = T;
    n_contrib[pix_id] = last_contributor;
    const size_t HW = (size_t)H * W;
#pragma unroll
    for (int ch = 0; ch < 3; ch++) out_color[ch * HW + pix_id] = C[ch] + T * bg_color[ch];
    if (F > 0) {
#pragma unroll
      for (int ch = 0; ch < F; ch++) out_lang[ch * HW + pix_id] = Fa[ch];
    }
    out_depth[pix_id] = D;
  }
  __syncthreads();
  if (tid == 0) tile_max_contrib[tile] = s_max;
}

template <int F>
static cudaError_t launch_rf(const uint2* ranges, const uint32_t* point_list, const float4* rec, int W, int H, int gx,
                             int gy, const float* bg, float* final_T, uint32_t* n_contrib, uint32_t* tile_max,
                             float* out_color, float* out_lang, float* out_depth, cudaStream_t st) {
  launch_pdl(render_fwd_kernel<F>, dim3(gx * gy), dim3(RF_THREADS), 0, st, ranges, point_list, rec, W, H, gx, bg, final_T, n_contrib, tile_max, out_color, out_lang, out_depth); count_launch();
  return cudaGetLastError();
}

cudaError_t launch_render_fwd(int F, const uint2* ranges, const uint32_t* point_list, const float4* rec, int W, int H,
                              int gx, int gy, const float* bg, float* final_T, uint32_t* n_contrib, uint32_t* tile_max,
                              float* out_color, float* out_lang, float* out_depth, cudaStream_t st) {
#define RF_CASE(N)                                                                                             \
  case N:                                                                                                      \
    return launch_rf<N>(ranges, point_list, rec, W, H, gx, gy, bg, final_T, n_contrib, tile_max, out_color, \
                        out_lang, out_depth, st);
  switch (F) {
    RF_CASE(0) RF_CASE(1) RF_CASE(2) RF_CASE(3) RF_CASE(4) RF_CASE(5) RF_CASE(6)
    default: return cudaErrorInvalidValue;
  }
#undef RF_CASE
}

}  // namespace fdls
