// render_bwd_shfl.cu -- the round-1 compositor backward, kept as the A/B partner of render_bwd.cu (FdlsBwdArgs.bwd_algo = 1):
// back-to-front re-traversal of every tile, producing per-Gaussian screen-space gradients.
// Replaces renderCUDA<3,F> backward (cuda_rasterizer/backward.cu:434-674).
//
// Per-pixel arithmetic (alpha recomputation, T /= (1-alpha), the accum_rec / last_* recurrences, the background
// term) follows the reference expression for expression.  What changes is how the per-pixel terms reach memory:
//
//  * reference: every contributing (pixel, Gaussian) pair issues 10+F scalar global atomicAdds (backward.cu:616-671).
//  * here: the 32 pixels of a warp look at the SAME Gaussian at the same time, so their 16 gradient terms are summed
//    inside the warp by a transposing butterfly (16 SHFL in total, not 16 x 5) that leaves term k in lane 2k; the
//    even lanes then issue one coalesced 64-byte vector of float atomics into the Gaussian's 64-byte gradient record.
//    => 32x fewer atomic operations reach L2 and they are sector-coalesced.
//  * warps in which no pixel is touched by the Gaussian skip it with one ballot.
//  * the traversal starts at the tile's deepest contributor (tile_max_contrib, written by render_fwd), not at the
//    end of the tile's range: the tail that every pixel would skip (backward.cu:578-580) is never staged.
//  * records are gathered with cp.async into a double-buffered shared ring exactly like the forward.
#include "common.cuh"

namespace fdls {

constexpr int RB_THREADS = TILE_PIX;
constexpr unsigned RB_FULL = 0xffffffffu;

__device__ __forceinline__ void rb_cp_async16(void* smem_dst, const void* gsrc) {
  uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void rb_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void rb_cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// Sum v[0..15] over the 32 lanes; on return lane 2k (k = 0..15) holds the total of term k.
__device__ __forceinline__ float warp_transpose_reduce16(float (&v)[16], int lane) {
  float w[8];
  {
    const bool hi = lane & 16;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float send = hi ? v[i] : v[i + 8];
      const float keep = hi ? v[i + 8] : v[i];
      w[i] = keep + __shfl_xor_sync(RB_FULL, send, 16);
    }
  }
  float x[4];
  {
    const bool hi = lane & 8;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float send = hi ? w[i] : w[i + 4];
      const float keep = hi ? w[i + 4] : w[i];
      x[i] = keep + __shfl_xor_sync(RB_FULL, send, 8);
    }
  }
  float y[2];
  {
    const bool hi = lane & 4;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const float send = hi ? x[i] : x[i + 2];
      const float keep = hi ? x[i + 2] : x[i];
      y[i] = keep + __shfl_xor_sync(RB_FULL, send, 4);
    }
  }
  float z;
  {
    const bool hi = lane & 2;
    const float send = hi ? y[0] : y[1];
    const float keep = hi ? y[1] : y[0];
    z = keep + __shfl_xor_sync(RB_FULL, send, 2);
  }
  z += __shfl_xor_sync(RB_FULL, z, 1);
  return z;
}

template <int F, bool DEPTH_GRAD>
__global__ void __launch_bounds__(RB_THREADS)
render_bwd_shfl_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                  const float4* __restrict__ rec, int W, int H, int gx, const float* __restrict__ bg_color,
                  const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                  const uint32_t* __restrict__ tile_max_contrib, const float* __restrict__ dL_dpixels,
                  const float* __restrict__ dL_dpixels_F, const float* __restrict__ dL_dpixel_depths,
                  float* __restrict__ grad_rec) {
  pdl_grid_sync();
  __shared__ float4 s_rec[2][4][RB_THREADS];
  __shared__ uint32_t s_id[2][RB_THREADS];
  __shared__ uint32_t s_strips[2][RB_THREADS];  // strip_summary().y of each staged Gaussian
  __shared__ uint8_t s_list[RB_THREADS / 32][RB_THREADS];

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = tid >> 5;
  const uint32_t strip_bit = 1u << warp;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const uint32_t tile = blockIdx.x;
  const uint32_t tx = tile % gx, ty = tile / gx;
  const uint2 pix = {tx * TILE + (tid & (TILE - 1)), ty * TILE + (tid >> 4)};
  const uint32_t pix_id = W * pix.y + pix.x;
  const float2 pixf = {(float)pix.x, (float)pix.y};
  const bool inside = pix.x < (uint32_t)W && pix.y < (uint32_t)H;
  const uint2 range = ranges[tile];
  // deepest position any pixel of the tile blended (<= range.y - range.x)
  const int M = (int)tile_max_contrib[tile];
  if (M == 0) return;
  const int rounds = (M + RB_THREADS - 1) / RB_THREADS;

  auto stage = [&](int round, int buf) {
    const int q = M - 1 - (round * RB_THREADS + tid);  // position inside the tile's range, back to front
    if (q >= 0) {
      const uint32_t id = point_list[range.x + q];
      s_id[buf][tid] = id;
      const float4* src = rec + (size_t)id * 4;
#pragma unroll
      for (int k = 0; k < 4; ++k) rb_cp_async16(&s_rec[buf][k][tid], src + k);
    }
    rb_cp_async_commit();
  };

  const float T_final = inside ? final_Ts[pix_id] : 0;
  float T = T_final;
  const int last_contributor = inside ? (int)n_contrib[pix_id] : 0;
  const size_t HW = (size_t)H * W;

  float accum_rec[3] = {0, 0, 0};
  float dL_dpixel[3] = {0, 0, 0};
  float dL_dpixel_depth = 0;
  float accum_depth_rec = 0;
  if (inside) {
#pragma unroll
    for (int i = 0; i < 3; i++) dL_dpixel[i] = dL_dpixels[i * HW + pix_id];
    if (DEPTH_GRAD) dL_dpixel_depth = dL_dpixel_depths[pix_id];
  }
  float last_alpha = 0;
  float last_color[3] = {0, 0, 0};
  float last_depth = 0;
  float accum_rec_F[F > 0 ? F : 1] = {0};
  float dL_dpixel_F[F > 0 ? F : 1] = {0};
  float last_lang[F > 0 ? F : 1] = {0};
  if (F > 0 && inside) {
#pragma unroll
    for (int i = 0; i < F; i++) dL_dpixel_F[i] = dL_dpixels_F[i * HW + pix_id];
  }
  // backward.cu:648-650 (loop-invariant; same accumulation order)
  float bg_dot_dpixel = 0;
#pragma unroll
  for (int i = 0; i < 3; i++) bg_dot_dpixel += bg_color[i] * dL_dpixel[i];

  const float ddelx_dx = 0.5 * W;
  const float ddely_dy = 0.5 * H;

  stage(0, 0);
  for (int i = 0; i < rounds; i++) {
    const int buf = i & 1;
    if (i + 1 < rounds) {
      stage(i + 1, buf ^ 1);
      rb_cp_async_wait<1>();
    } else {
      rb_cp_async_wait<0>();
    }
    {
      const float4 r0 = s_rec[buf][0][tid], r1 = s_rec[buf][1][tid];
      s_strips[buf][tid] = __float_as_uint(
          strip_summary(r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, (float)(tx * TILE), (float)(ty * TILE)).y);
    }
    __syncthreads();
    const int cnt = min(RB_THREADS, M - i * RB_THREADS);
    // per-warp compaction of the batch to the Gaussians that can reach this warp's strip (see render_fwd.cu)
    int n_hit = 0;
#pragma unroll
    for (int k = 0; k < RB_THREADS / 32; ++k) {
      const int jj = k * 32 + lane;
      const bool hit = jj < cnt && (s_strips[buf][jj] & strip_bit);
      const uint32_t m = __ballot_sync(RB_FULL, hit);
      if (hit) s_list[warp][n_hit + __popc(m & lt_mask)] = (uint8_t)jj;
      n_hit += __popc(m);
    }
    __syncwarp();
    for (int h = 0; h < n_hit; h++) {
      const int j = s_list[warp][h];
      const int q = M - 1 - (i * RB_THREADS + j);  // == the reference's `contributor` after its decrement
      const float4 a = s_rec[buf][0][j];           // x, y, conic.x, conic.y
      const float4 b = s_rec[buf][1][j];           // conic.z, opacity, depth, r
      const float2 d = {a.x - pixf.x, a.y - pixf.y};
      const float power = -0.5f * (a.z * d.x * d.x + b.x * d.y * d.y) - a.w * d.x * d.y;
      bool act = q < last_contributor && !(power > 0.0f);
      float G = 0, alpha = 0;
      if (act) {
        G = exp(power);
        alpha = min(0.99f, b.y * G);
        act = !(alpha < 1.0f / 255.0f);
      }
      if (!__any_sync(RB_FULL, act)) continue;

      float v[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) v[k] = 0.f;
      if (act) {
        T = T / (1.f - alpha);
        const float dchannel_dcolor = alpha * T;
        float dL_dalpha = 0.0f;
        const float4 c4 = s_rec[buf][2][j];  // g, b, l0, l1
        const float col[3] = {b.w, c4.x, c4.y};
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
          const float c = col[ch];
          accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
          last_color[ch] = c;
          const float dL_dchannel = dL_dpixel[ch];
          dL_dalpha += (c - accum_rec[ch]) * dL_dchannel;
          v[G_RGB + ch] = dchannel_dcolor * dL_dchannel;
        }
        if (F > 0) {
          const float4 e4 = s_rec[buf][3][j];
          const float l[6] = {c4.z, c4.w, e4.x, e4.y, e4.z, e4.w};
#pragma unroll
          for (int ch = 0; ch < F; ch++) {
            const float f = l[ch];
            accum_rec_F[ch] = last_alpha * last_lang[ch] + (1.f - last_alpha) * accum_rec_F[ch];
            last_lang[ch] = f;
            const float dL_dchannel_F = dL_dpixel_F[ch];
            dL_dalpha += (f - accum_rec_F[ch]) * dL_dchannel_F;
            v[g_lang_slot(ch)] = dchannel_dcolor * dL_dchannel_F;
          }
        }
        if (DEPTH_GRAD) {
          const float c_d = b.z;
          accum_depth_rec = last_alpha * last_depth + (1.f - last_alpha) * accum_depth_rec;
          last_depth = c_d;
          dL_dalpha += (c_d - accum_depth_rec) * dL_dpixel_depth;
          v[G_DEPTH] = dchannel_dcolor * dL_dpixel_depth;
        }
        dL_dalpha *= T;
        last_alpha = alpha;
        dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot_dpixel;

        const float dL_dG = b.y * dL_dalpha;
        const float gdx = G * d.x;
        const float gdy = G * d.y;
        const float dG_ddelx = -gdx * a.z - gdy * a.w;
        const float dG_ddely = -gdy * b.x - gdx * a.w;
        v[G_MX] = dL_dG * dG_ddelx * ddelx_dx;
        v[G_MY] = dL_dG * dG_ddely * ddely_dy;
        v[G_CA] = -0.5f * gdx * d.x * dL_dG;
        v[G_CB] = -0.5f * gdx * d.y * dL_dG;
        v[G_CC] = -0.5f * gdy * d.y * dL_dG;
        v[G_OP] = G * dL_dalpha;
      }
      const float total = warp_transpose_reduce16(v, lane);
      const int term = lane >> 1;
      if ((lane & 1) == 0 && ((g_used_slots(F, DEPTH_GRAD) >> term) & 1u)) atomicAdd(grad_rec + (size_t)s_id[buf][j] * 16 + term, total);
    }
    __syncthreads();
  }
}

template <int F>
static cudaError_t launch_rb_shfl(bool depth_grad, const uint2* ranges, const uint32_t* point_list, const float4* rec, int W,
                             int H, int gx, int gy, const float* bg, const float* final_T, const uint32_t* n_contrib,
                             const uint32_t* tile_max, const float* dpix, const float* dpix_F, const float* dpix_D,
                             float* grad_rec, cudaStream_t st) {
  count_launch();
  if (depth_grad)
    launch_pdl(render_bwd_shfl_kernel<F, true>, dim3(gx * gy), dim3(RB_THREADS), 0, st, ranges, point_list, rec, W, H, gx, bg, final_T, n_contrib, tile_max, dpix, dpix_F, dpix_D, grad_rec);
  else
    launch_pdl(render_bwd_shfl_kernel<F, false>, dim3(gx * gy), dim3(RB_THREADS), 0, st, ranges, point_list, rec, W, H, gx, bg, final_T, n_contrib, tile_max, dpix, dpix_F, dpix_D, grad_rec);
  return cudaGetLastError();
}

cudaError_t launch_render_bwd_shfl(int F, const uint2* ranges, const uint32_t* point_list, const float4* rec, int W, int H,
                              int gx, int gy, const float* bg, const float* final_T, const uint32_t* n_contrib,
                              const uint32_t* tile_max, const float* dpix, const float* dpix_F, const float* dpix_D,
                              float* grad_rec, cudaStream_t st) {
  const bool dg = dpix_D != nullptr;
#define RB_CASE(N)                                                                                                \
  case N:                                                                                                         \
    return launch_rb_shfl<N>(dg, ranges, point_list, rec, W, H, gx, gy, bg, final_T, n_contrib, tile_max, dpix, dpix_F, \
                        dpix_D, grad_rec, st);
  switch (F) {
    RB_CASE(0) RB_CASE(1) RB_CASE(2) RB_CASE(3) RB_CASE(4) RB_CASE(5) RB_CASE(6)
    default: return cudaErrorInvalidValue;
  }
#undef RB_CASE
}

}  // namespace fdls
