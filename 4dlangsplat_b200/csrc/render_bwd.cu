// render_bwd.cu -- back-to-front re-traversal of every tile, producing per-Gaussian screen-space gradients.
// Replaces renderCUDA<3,F> backward (cuda_rasterizer/backward.cu:434-674).
//
// The traversal is the reference's: one pixel per thread, back to front from the tile's deepest contributor, alpha
// recomputed with the same expressions and the same two tests (power > 0, alpha < 1/255), T divided back out.
// What is new is how the per-(pixel, Gaussian) terms become per-Gaussian sums.
//
//  * reference: every contributing pair issues 10+F scalar global atomicAdds (backward.cu:616-671).
//  * round 1 of this repo (render_bwd_shfl.cu): the 16 terms of a pair were summed over the warp with a transposing
//    butterfly -- ~65 of the ~190 warp-instructions a blended (warp, Gaussian) step cost.
//  * here the sums over pixels are two small GEMMs per warp, run on the tensor cores:
//      channel block:   dL_dchannel[g][ch] = sum_pix w[g][pix]  * dL_dpixel[pix][ch]      w  = alpha * T
//      geometry block:  m[g][k]            = sum_pix u[g][pix]  * phi[pix][k]             u  = G * dL_dalpha
//    with phi = {1, x, y, x^2, xy, y^2} of the pixel's coordinates relative to the centre of the warp's 8x4 block.  Because
//    d = (gauss - centre) - (pixel - centre), the five geometric gradients of a Gaussian (mean2D x/y, conic a/b/c)
//    and dL_dopacity are closed-form combinations of these six moments (flush_group below), so a pair contributes
//    only TWO numbers (w, u) instead of 10+F products.  Per pair the 3+F vector recurrences of
//    accum_rec[ch] (backward.cu:598-611) collapse into ONE scalar recurrence of A = sum_ch accum_rec[ch]*dL_dpixel[ch],
//    since dL_dpixel is constant along a pixel's traversal.
//    Each warp parks the (w, u) rows of the Gaussians it blends in shared memory (one row of 32 pixels per Gaussian);
//    every 16 rows it issues mma.sync.m16n8k8 TF32 with a 3-term split of the operands
//    (w = hi + lo, dL = hi + lo; u split EXACTLY in three, phi exact in TF32: |x| <= 3.5, |y| <= 1.5), fp32
//    accumulation, i.e. fp32-grade sums (each warp owns an 8x4-pixel block, not a 16x2 strip, to keep |phi| small:
//    the closed forms subtract moments, and their cancellation grows with the block's extent);
//    the quad that holds a Gaussian's row of the result emits it as two 32-byte vector reductions
//    (red.global.add.v2.f32 x 4 lanes) into the Gaussian's 64-byte gradient record.
//    => per blended step: ~20 FP32 instructions of recurrence + 2 shared stores + 1/16 of a flush, against 75 + 65.
//  * tcgen05 is not used on purpose: the M dimension is a per-WARP list of 16 Gaussians and the operand changes every
//    16 steps; a CTA-wide UMMA would need the rows every other warp skipped zero-filled (10x the stores).
//  * warps in which no pixel is touched by the Gaussian skip it with one ballot; the traversal starts at the tile's
//    deepest contributor (tile_max_contrib, written by render_fwd); records are gathered with cp.async into a
//    double-buffered shared ring exactly like the forward.
#include <cstddef>
#include "common.cuh"

namespace fdls {

constexpr int RB_THREADS = TILE_PIX;
constexpr int RB_WARPS = RB_THREADS / 32;
constexpr int RB_BATCH = 128;   // Gaussians staged per round (two threads gather one 64-byte record)
constexpr unsigned RB_FULL = 0xffffffffu;
constexpr int RB_GROUP = 16;    // Gaussians per MMA group (the M of m16n8k8)
constexpr int RB_ASTRIDE = 32;  // floats per parked row (32 pixels); 16-byte groups XOR-swizzled by the row, see rb_col
constexpr int RB_BSTRIDE = 36;  // floats per channel row of the B operand: 32 pixels + 4 => conflict-free 128-bit loads
constexpr int RB_STASHED = 0x100;

// Shared memory of one CTA.  NCH = rows of the B operand (8, or 10 when language channels 4/5 exist).
template <int NCH>
struct RbSmem {
  float4 rec[2][4][RB_BATCH];
  uint32_t id[2][RB_BATCH];
  uint32_t blocks[2][RB_BATCH];  // block_summary() of each staged Gaussian
  uint8_t list[RB_WARPS][RB_BATCH];
  float w[RB_WARPS][RB_GROUP * RB_ASTRIDE];
  float u[RB_WARPS][RB_GROUP * RB_ASTRIDE];
  float b_hi[RB_WARPS][NCH * RB_BSTRIDE];  // dL_dpixel per channel and pixel of the warp's block, TF32 high part
  float b_lo[RB_WARPS][NCH * RB_BSTRIDE];  //   and the remainder
  float4 st_a[RB_WARPS][RB_GROUP];  // parameters of rows carried over a batch boundary: x, y, conic a, b
  float2 st_b[RB_WARPS][RB_GROUP];  //   conic c, opacity
  uint32_t st_id[RB_WARPS][RB_GROUP];
};

__device__ __forceinline__ void rb_cp_async16(void* smem_dst, const void* gsrc) {
  uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void rb_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void rb_cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// Word of row g that holds pixel (lane) s: the eight 16-byte groups of a row are XOR-permuted by the row index, so that
// the 128-bit fragment loads of a flush (8 lanes of a quarter-warp: rows g, g+1 x groups 2*tig + h) hit 8 different
// bank groups, and the per-step stores (32 lanes, one row) stay a permutation of the banks.  No padding: 3 CTAs / SM.
__device__ __forceinline__ int rb_col(int g, int s) { return s ^ ((g & 7) << 2); }

// TF32 operand splits.  The tensor core reads only the upper 19 bits of an operand register, so `x & 0xffffe000` IS
// the TF32 value the hardware would see, and x minus it is exact in fp32.  (cvt.rna.tf32.f32 has no single-instruction
// form on sm_100a -- ptxas expands it to ~5 integer instructions -- so rounding the split cost 4x more than the MMAs.)
//   2-way: x = hi + lo, lo passed as is (the hardware truncates it to 11 bits: error < 2^-20 |x|)
//   3-way: x = h1 + h2 + h3 EXACTLY (11 + 11 + 2 significant bits)
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(x) & 0xffffe000u;
  lo = __float_as_uint(x - __uint_as_float(hi));
}
__device__ __forceinline__ void split3_tf32(float x, uint32_t& h1, uint32_t& h2, uint32_t& h3) {
  h1 = __float_as_uint(x) & 0xffffe000u;
  const float r1 = x - __uint_as_float(h1);
  h2 = __float_as_uint(r1) & 0xffffe000u;
  h3 = __float_as_uint(r1 - __uint_as_float(h2));
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

template <int F, bool DEPTH_GRAD>
__global__ void __launch_bounds__(RB_THREADS, F > 4 ? 2 : 3)
render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                  const float4* __restrict__ rec, int W, int H, int gx, const float* __restrict__ bg_color,
                  const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                  const uint32_t* __restrict__ tile_max_contrib, const float* __restrict__ dL_dpixels,
                  const float* __restrict__ dL_dpixels_F, const float* __restrict__ dL_dpixel_depths,
                  float* __restrict__ grad_rec, const uint32_t* __restrict__ tile_order) {
  pdl_grid_sync();
  constexpr bool TWO_BLOCKS = F > 4;  // language channels 4, 5 need a second 8-column block
  constexpr int NCH = TWO_BLOCKS ? 10 : 8;
  constexpr int N_CH = F > 0 ? 4 + (F < 4 ? F : 4) : (DEPTH_GRAD ? 4 : 3);  // used columns of the first block
  extern __shared__ __align__(16) unsigned char rb_smem_raw[];
  RbSmem<NCH>& sm = *reinterpret_cast<RbSmem<NCH>*>(rb_smem_raw);

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = tid >> 5;
  const int gid = lane >> 2, tig = lane & 3;  // mma fragment coordinates
  const uint32_t block_bit = 1u << warp;      // this warp's 8x4-pixel block of the tile: bx = warp & 1, by = warp >> 1
  const int bx0 = (warp & 1) * 8, by0 = (warp >> 1) * 4;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const uint32_t tile = tile_order != nullptr ? tile_order[blockIdx.x] : blockIdx.x;  // deepest tiles first
  const uint32_t tx = tile % gx, ty = tile / gx;
  const uint2 pix = {tx * TILE + bx0 + (lane & 7), ty * TILE + by0 + (lane >> 3)};
  const uint32_t pix_id = W * pix.y + pix.x;
  const float2 pixf = {(float)pix.x, (float)pix.y};
  const bool inside = pix.x < (uint32_t)W && pix.y < (uint32_t)H;
  const uint2 range = ranges[tile];
  // deepest position any pixel of the tile blended (<= range.y - range.x)
  const int M = (int)tile_max_contrib[tile];
  if (M == 0) return;
  const int rounds = (M + RB_BATCH - 1) / RB_BATCH;
  const size_t HW = (size_t)H * W;

  // two threads gather one record: thread t copies slots 2*(t&1), 2*(t&1)+1 of the Gaussian at batch position t>>1
  auto stage = [&](int round, int buf) {
    const int slot = tid >> 1, half = tid & 1;
    const int q = M - 1 - (round * RB_BATCH + slot);  // position inside the tile's range, back to front
    if (q >= 0) {
      const uint32_t id = point_list[range.x + q];
      if (half == 0) sm.id[buf][slot] = id;
      const float4* src = rec + (size_t)id * 4 + 2 * half;
      rb_cp_async16(&sm.rec[buf][2 * half][slot], src);
      rb_cp_async16(&sm.rec[buf][2 * half + 1][slot], src + 1);
    }
    rb_cp_async_commit();
  };
  stage(0, 0);

  // ---- per-pixel state
  const float T_final = inside ? final_Ts[pix_id] : 0;
  float T = T_final;
  const int last_contributor = inside ? (int)n_contrib[pix_id] : 0;
  float dL_dpixel[3] = {0, 0, 0};
  float dL_dpixel_depth = 0;
  float dL_dpixel_F[F > 0 ? F : 1] = {0};
  if (inside) {
#pragma unroll
    for (int i = 0; i < 3; i++) dL_dpixel[i] = dL_dpixels[i * HW + pix_id];
    if (DEPTH_GRAD) dL_dpixel_depth = dL_dpixel_depths[pix_id];
#pragma unroll
    for (int i = 0; i < F; i++) dL_dpixel_F[i] = dL_dpixels_F[i * HW + pix_id];
  }
  float bg_dot_dpixel = 0;  // backward.cu:648-650
#pragma unroll
  for (int i = 0; i < 3; i++) bg_dot_dpixel += bg_color[i] * dL_dpixel[i];
  const float bg_term = -T_final * bg_dot_dpixel;
  float A = 0;  // sum over channels of accum_rec[ch] * dL_dpixel[ch] (backward.cu:598-611 folded into one recurrence)

  // ---- B operand of the channel block (constant for the whole tile): row = record slot of the channel
  // (G_RGB.., G_DEPTH, G_LANG.., then language channels 4, 5), column = pixel (lane), split hi + lo for the 3-term
  // TF32 product.  Written once by the lane that owns the pixel; read as fragments at every flush.
  {
    float* bh = sm.b_hi[warp];
    float* bl = sm.b_lo[warp];
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      float v = 0.f;
      if (c < 3) v = dL_dpixel[c];
      else if (c == 3) v = DEPTH_GRAD ? dL_dpixel_depth : 0.f;
      else if (c < 8) v = (c - 4 < F) ? dL_dpixel_F[c - 4 < F ? c - 4 : 0] : 0.f;
      else v = (c - 4 < F) ? dL_dpixel_F[c - 4 < F ? c - 4 : 0] : 0.f;
      uint32_t hi, lo;
      split_tf32(v, hi, lo);
      bh[c * RB_BSTRIDE + lane] = __uint_as_float(hi);
      bl[c * RB_BSTRIDE + lane] = __uint_as_float(lo);
    }
  }
  // geometry block: phi_n(pixel) for n = gid, as c0 + c1 * x + c2 * x^2 with x = e - 3.5 (the fragment element index e
  // IS the pixel's column in the 8x4 block) and y = tig - 1.5 (tig IS its row): {1, x, y, x^2, x y, y^2, 0, 0}
  const float yq = (float)tig - 1.5f;
  const float phi_c0 = gid == 0 ? 1.f : gid == 2 ? yq : gid == 5 ? yq * yq : 0.f;
  const float phi_c1 = gid == 1 ? 1.f : gid == 4 ? yq : 0.f;
  const float phi_c2 = gid == 3 ? 1.f : 0.f;
  const float cx = (float)(tx * TILE + bx0) + 3.5f, cy = (float)(ty * TILE + by0) + 1.5f;
  const float half_W = 0.5f * W, half_H = 0.5f * H;  // ddelx_dx, ddely_dy (backward.cu:507-508)

  // byte offset of this lane inside a parked row, kept opaque so that it stays in a register (under the 80-register
  // cap the compiler otherwise re-derives it from %tid every step)
  uint32_t lane4, park_w;  // park_w: shared-window address of this warp's first parked row (sm.u is sizeof(sm.w) above)
  // (a plain mov is transparent to ptxas, which then rebuilds lane * 4 from %tid -- S2R + 2 ALU -- in EVERY step; the
  // result of a volatile shuffle cannot be rematerialised)
  asm volatile("shfl.sync.idx.b32 %0, %1, %2, 0x1f, 0xffffffff;" : "=r"(lane4) : "r"(lane * 4), "r"(lane));
  asm volatile("mov.u32 %0, %1;" : "=r"(park_w) : "r"((uint32_t)__cvta_generic_to_shared(sm.w[warp])));
  static_assert(offsetof(RbSmem<NCH>, u) - offsetof(RbSmem<NCH>, w) == sizeof(float) * RB_WARPS * RB_GROUP * RB_ASTRIDE,
                "sm.u must follow sm.w");
  int n_pend = 0;  // rows parked in sm.w / sm.u of this warp (warp-uniform)
  int my_j = 0;    // lane r < n_pend: where row r's Gaussian lives (index into the current batch, or RB_STASHED)
  float* const row_w = sm.w[warp];
  float* const row_u = sm.u[warp];

  // Reduce the parked rows [0, n_rows) over this warp's 32 pixels and add the result to the gradient records.
  auto flush_group = [&](int n_rows, int buf) {
    float d1[4] = {0, 0, 0, 0}, d2[4] = {0, 0, 0, 0}, dm[4] = {0, 0, 0, 0};
    // rows gid and gid + 8 share gid & 7, hence the same swizzle
    const int c0 = rb_col(gid, tig * 8), c1 = rb_col(gid, tig * 8 + 4);
    {
      const float* rw0 = row_w + gid * RB_ASTRIDE;
      const float* rw1 = row_w + (gid + 8) * RB_ASTRIDE;
      const float4 p0 = *reinterpret_cast<const float4*>(rw0 + c0), p1 = *reinterpret_cast<const float4*>(rw0 + c1);
      const float4 q0 = *reinterpret_cast<const float4*>(rw1 + c0), q1 = *reinterpret_cast<const float4*>(rw1 + c1);
      const float lo_row[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
      const float hi_row[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
      const float* bh = sm.b_hi[warp] + gid * RB_BSTRIDE + tig * 8;
      const float* bl = sm.b_lo[warp] + gid * RB_BSTRIDE + tig * 8;
      const float4 h0 = *reinterpret_cast<const float4*>(bh), h1 = *reinterpret_cast<const float4*>(bh + 4);
      const float4 l0 = *reinterpret_cast<const float4*>(bl), l1 = *reinterpret_cast<const float4*>(bl + 4);
      const float bhi[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
      const float blo[8] = {l0.x, l0.y, l0.z, l0.w, l1.x, l1.y, l1.z, l1.w};
      float b2hi[8] = {0, 0, 0, 0, 0, 0, 0, 0}, b2lo[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      if (TWO_BLOCKS && gid >= 6) {  // columns 6, 7 of the second block = language channels 4, 5 (B rows 8, 9)
        const float* ch = sm.b_hi[warp] + (gid + 2) * RB_BSTRIDE + tig * 8;
        const float* cl = sm.b_lo[warp] + (gid + 2) * RB_BSTRIDE + tig * 8;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          b2hi[e] = ch[e];
          b2lo[e] = cl[e];
        }
      }
#pragma unroll
      for (int kb = 0; kb < 4; ++kb) {
        uint32_t ah[4], al[4];
        split_tf32(lo_row[2 * kb], ah[0], al[0]);
        split_tf32(hi_row[2 * kb], ah[1], al[1]);
        split_tf32(lo_row[2 * kb + 1], ah[2], al[2]);
        split_tf32(hi_row[2 * kb + 1], ah[3], al[3]);
        const uint32_t bh0 = __float_as_uint(bhi[2 * kb]), bh1 = __float_as_uint(bhi[2 * kb + 1]);
        mma_tf32(d1, al, bh0, bh1);
        mma_tf32(d1, ah, __float_as_uint(blo[2 * kb]), __float_as_uint(blo[2 * kb + 1]));
        mma_tf32(d1, ah, bh0, bh1);
        if (TWO_BLOCKS) {
          const uint32_t ch0 = __float_as_uint(b2hi[2 * kb]), ch1 = __float_as_uint(b2hi[2 * kb + 1]);
          mma_tf32(d2, al, ch0, ch1);
          mma_tf32(d2, ah, __float_as_uint(b2lo[2 * kb]), __float_as_uint(b2lo[2 * kb + 1]));
          mma_tf32(d2, ah, ch0, ch1);
        }
      }
    }
    {
      const float* ru0 = row_u + gid * RB_ASTRIDE;
      const float* ru1 = row_u + (gid + 8) * RB_ASTRIDE;
      const float4 p0 = *reinterpret_cast<const float4*>(ru0 + c0), p1 = *reinterpret_cast<const float4*>(ru0 + c1);
      const float4 q0 = *reinterpret_cast<const float4*>(ru1 + c0), q1 = *reinterpret_cast<const float4*>(ru1 + c1);
      const float lo_row[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
      const float hi_row[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
      for (int kb = 0; kb < 4; ++kb) {
        // u = h1 + h2 + h3 EXACTLY (11 + 11 + 2 significant bits) and phi is exact, so every product is exact and
        // only the fp32 accumulation rounds: the closed forms below subtract moments, their operands must not be noisy
        uint32_t a1[4], a2[4], a3[4];
        split3_tf32(lo_row[2 * kb], a1[0], a2[0], a3[0]);
        split3_tf32(hi_row[2 * kb], a1[1], a2[1], a3[1]);
        split3_tf32(lo_row[2 * kb + 1], a1[2], a2[2], a3[2]);
        split3_tf32(hi_row[2 * kb + 1], a1[3], a2[3], a3[3]);
        const float x0 = (float)(2 * kb) - 3.5f, x1 = (float)(2 * kb + 1) - 3.5f;
        const uint32_t f0 = __float_as_uint(phi_c0 + phi_c1 * x0 + phi_c2 * (x0 * x0));
        const uint32_t f1 = __float_as_uint(phi_c0 + phi_c1 * x1 + phi_c2 * (x1 * x1));
        mma_tf32(dm, a3, f0, f1);
        mma_tf32(dm, a2, f0, f1);
        mma_tf32(dm, a1, f0, f1);
      }
    }
    // d1[0..1] = channel columns 2*tig, 2*tig+1 of row gid, d1[2..3] of row gid+8; dm likewise = the six moments
    // {m00, m10 | m01, m20 | m11, m02 | -, -} spread over the quad.
    const int qbase = lane & ~3;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int r = gid + 8 * half;
      const int jf = __shfl_sync(RB_FULL, my_j, r);
      const float e0 = dm[2 * half], e1 = dm[2 * half + 1];
      const float m00 = __shfl_sync(RB_FULL, e0, qbase), m10 = __shfl_sync(RB_FULL, e1, qbase);
      const float m01 = __shfl_sync(RB_FULL, e0, qbase + 1), m20 = __shfl_sync(RB_FULL, e1, qbase + 1);
      const float m11 = __shfl_sync(RB_FULL, e0, qbase + 2), m02 = __shfl_sync(RB_FULL, e1, qbase + 2);
      if (r < n_rows) {
        float4 pa;
        float2 pb;
        uint32_t id;
        if (jf & RB_STASHED) {
          pa = sm.st_a[warp][r];
          pb = sm.st_b[warp][r];
          id = sm.st_id[warp][r];
        } else {
          pa = sm.rec[buf][0][jf];
          const float4 t = sm.rec[buf][1][jf];
          pb = make_float2(t.x, t.y);
          id = sm.id[buf][jf];
        }
        // d = g - p with g, p taken relative to the block centre; u-weighted sums of 1, dx, dy, dx^2, dx dy, dy^2
        const float gxr = pa.x - cx, gyr = pa.y - cy;
        const float Sx = gxr * m00 - m10, Sy = gyr * m00 - m01;
        const float Sxx = gxr * (Sx - m10) + m20;
        const float Sxy = gxr * Sy - gyr * m10 + m11;
        const float Syy = gyr * (Sy - m01) + m02;
        const float op = pb.y;
        float2 geo;
        if (tig == 0) geo = make_float2(-op * (pa.z * Sx + pa.w * Sy) * half_W, -op * (pb.x * Sy + pa.w * Sx) * half_H);
        else if (tig == 1) geo = make_float2(-0.5f * op * Sxx, -0.5f * op * Sxy);
        else geo = make_float2(-0.5f * op * Syy, m00);
        float* g = grad_rec + (size_t)id * 16;
        if (2 * tig < N_CH) atomicAdd(reinterpret_cast<float2*>(g + 2 * tig), make_float2(d1[2 * half], d1[2 * half + 1]));
        if (tig < 3) atomicAdd(reinterpret_cast<float2*>(g + G_MX + 2 * tig), geo);
        else if (TWO_BLOCKS) atomicAdd(reinterpret_cast<float2*>(g + G_LANG_HI), make_float2(d2[2 * half], d2[2 * half + 1]));
      }
    }
  };

  for (int i = 0; i < rounds; i++) {
    const int buf = i & 1;
    if (i + 1 < rounds) {
      stage(i + 1, buf ^ 1);
      rb_cp_async_wait<1>();
    } else {
      rb_cp_async_wait<0>();
    }
    __syncwarp();  // the two lanes that gathered a record see each other's half
    const int cnt = min(RB_BATCH, M - i * RB_BATCH);
    {
      const int slot = tid >> 1;
      uint32_t bits = 0;
      if (slot < cnt) {
        const float4 r0 = sm.rec[buf][0][slot], r1 = sm.rec[buf][1][slot];
        bits = block_summary(r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, (float)(tx * TILE), (float)(ty * TILE), tid & 1);
      }
      bits |= __shfl_xor_sync(RB_FULL, bits, 1);  // lane pair = block columns bx 0 and 1
      if ((tid & 1) == 0) sm.blocks[buf][slot] = bits;
    }
    __syncthreads();
    // per-warp compaction of the batch to the Gaussians that can reach this warp's block (as in render_fwd.cu)
    int n_hit = 0;
#pragma unroll
    for (int k = 0; k < RB_BATCH / 32; ++k) {
      const int jj = k * 32 + lane;
      const bool hit = jj < cnt && (sm.blocks[buf][jj] & block_bit);
      const uint32_t m = __ballot_sync(RB_FULL, hit);
      if (hit) sm.list[warp][n_hit + __popc(m & lt_mask)] = (uint8_t)jj;
      n_hit += __popc(m);
    }
    __syncwarp();
    // `contributor` of the reference after its decrement is q = M - 1 - (i * RB_BATCH + j); q < last_contributor
    const int j_min = M - 1 - i * RB_BATCH - last_contributor;  // active iff j > j_min
    for (int h = 0; h < n_hit; h++) {
      const int j = sm.list[warp][h];
      const float4 a = sm.rec[buf][0][j];  // x, y, conic.x, conic.y
      const float4 b = sm.rec[buf][1][j];  // conic.z, opacity, depth, r
      const float2 d = {a.x - pixf.x, a.y - pixf.y};
      const float power = -0.5f * (a.z * d.x * d.x + b.x * d.y * d.y) - a.w * d.x * d.y;
      bool act = j > j_min && !(power > 0.0f);
      float G = 0, alpha = 0;
      if (act) {
        G = exp(power);
        alpha = min(0.99f, b.y * G);
        act = !(alpha < 1.0f / 255.0f);
      }
      if (!__any_sync(RB_FULL, act)) continue;

      float w = 0.f, u = 0.f;
      if (act) {
        const float rinv = rcp_approx(1.f - alpha);
        T = T * rinv;   // backward.cu:589
        w = alpha * T;  // dchannel_dcolor
        const float4 c4 = sm.rec[buf][2][j];  // g, b, l0, l1
        float S = b.w * dL_dpixel[0] + c4.x * dL_dpixel[1] + c4.y * dL_dpixel[2];
        if (DEPTH_GRAD) S += b.z * dL_dpixel_depth;
        if (F > 0) {
          const float l01[2] = {c4.z, c4.w};
#pragma unroll
          for (int ch = 0; ch < (F < 2 ? F : 2); ch++) S += l01[ch] * dL_dpixel_F[ch];
          if (F > 2) {
            const float4 e4 = sm.rec[buf][3][j];
            const float l25[4] = {e4.x, e4.y, e4.z, e4.w};
#pragma unroll
            for (int ch = 2; ch < F; ch++) S += l25[ch - 2] * dL_dpixel_F[ch];
          }
        }
        // dL_dalpha = T * sum_ch (c[ch] - accum_rec[ch]) * dL_dpixel[ch] + (-T_final / (1 - alpha)) * bg . dL_dpixel
        const float diff = S - A;
        const float dL_dalpha = diff * T + bg_term * rinv;
        A += alpha * diff;  // accum_rec <- alpha * c + (1 - alpha) * accum_rec, dotted with dL_dpixel
        u = G * dL_dalpha;
      }
      // row n_pend, word rb_col(n_pend, lane)
      const uint32_t addr = park_w + (uint32_t)n_pend * (RB_ASTRIDE * 4) + (lane4 ^ ((uint32_t)(n_pend & 7) << 4));
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(w) : "memory");
      asm volatile("st.shared.f32 [%0+%2], %1;" ::"r"(addr), "f"(u), "n"(sizeof(float) * RB_WARPS * RB_GROUP * RB_ASTRIDE)
                   : "memory");
      if (lane4 == (uint32_t)n_pend * 4) my_j = j;
      if (++n_pend == RB_GROUP) {
        __syncwarp();
        flush_group(RB_GROUP, buf);
        n_pend = 0;
        __syncwarp();
      }
    }
    // rows still parked belong to Gaussians of this batch: keep their parameters before the ring slot is reused
    if (n_pend > 0 && i + 1 < rounds) {
      if (lane < n_pend && !(my_j & RB_STASHED)) {
        sm.st_a[warp][lane] = sm.rec[buf][0][my_j];
        const float4 t = sm.rec[buf][1][my_j];
        sm.st_b[warp][lane] = make_float2(t.x, t.y);
        sm.st_id[warp][lane] = sm.id[buf][my_j];
        my_j = RB_STASHED;
      }
    }
    if (i + 1 == rounds && n_pend > 0) {
      __syncwarp();
      flush_group(n_pend, buf);
    }
    __syncthreads();
  }
}

// First kernel of a backward: zeroes the gradient records and, in block 0, orders the tiles by the depth of their
// traversal (tile_max_contrib, i.e. by their cost), deepest first.  The hardware hands out CTAs in blockIdx order, so
// this is longest-processing-time-first scheduling: with 2040 tiles on 444 CTA slots the unordered launch lost ~10 % of
// the kernel's time to the last wave (SMs idle while others finish a deep tile that started late).  64 buckets of
// 4 * log2(M + 1); the order inside a bucket is arbitrary.
constexpr int RB_ORDER_BUCKETS = 64;
__global__ void __launch_bounds__(256)
bwd_prepare_kernel(float4* __restrict__ grad_rec, size_t n_vec, const uint32_t* __restrict__ tile_max, int tiles,
                   uint32_t* __restrict__ tile_order) {
  pdl_grid_sync();
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  const bool order = tile_order != nullptr;  // then block 0 only orders, the others zero
  if (!order || blockIdx.x != 0) {
    const size_t nb = gridDim.x - (order ? 1 : 0), bi = blockIdx.x - (order ? 1 : 0);
    for (size_t i = bi * blockDim.x + threadIdx.x; i < n_vec; i += nb * blockDim.x) grad_rec[i] = z;
    return;
  }
  __shared__ uint32_t s_count[RB_ORDER_BUCKETS];
  auto bucket = [](uint32_t M) {
    const int b = (int)(4.f * log2f((float)M + 1.f));
    return RB_ORDER_BUCKETS - 1 - (b < RB_ORDER_BUCKETS - 1 ? b : RB_ORDER_BUCKETS - 1);  // deepest -> bucket 0
  };
  if (threadIdx.x < RB_ORDER_BUCKETS) s_count[threadIdx.x] = 0;
  __syncthreads();
  for (int t = threadIdx.x; t < tiles; t += blockDim.x) atomicAdd(&s_count[bucket(tile_max[t])], 1u);
  __syncthreads();
  if (threadIdx.x < 32) {  // exclusive scan of the 64 counts
    const uint32_t a = s_count[2 * threadIdx.x], b = s_count[2 * threadIdx.x + 1];
    uint32_t incl = a + b;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)threadIdx.x >= o) incl += v;
    }
    s_count[2 * threadIdx.x] = incl - a - b;
    s_count[2 * threadIdx.x + 1] = incl - b;
  }
  __syncthreads();
  for (int t = threadIdx.x; t < tiles; t += blockDim.x) tile_order[atomicAdd(&s_count[bucket(tile_max[t])], 1u)] = (uint32_t)t;
}

cudaError_t launch_bwd_prepare(float* grad_rec, size_t n_floats, const uint32_t* tile_max, int tiles, uint32_t* tile_order,
                               cudaStream_t st) {
  count_launch();
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  launch_pdl(bwd_prepare_kernel, dim3(sms * 4), dim3(256), 0, st, reinterpret_cast<float4*>(grad_rec), n_floats / 4, tile_max, tiles, tile_order);
  return cudaGetLastError();
}

template <int F>
static cudaError_t launch_rb(bool depth_grad, const uint2* ranges, const uint32_t* point_list, const float4* rec, int W,
                             int H, int gx, int gy, const float* bg, const float* final_T, const uint32_t* n_contrib,
                             const uint32_t* tile_max, const float* dpix, const float* dpix_F, const float* dpix_D,
                             float* grad_rec, const uint32_t* tile_order, cudaStream_t st) {
  static SmemOptIn opt_d, opt_n;
  constexpr size_t smem = sizeof(RbSmem<(F > 4 ? 10 : 8)>);
  count_launch();
  if (depth_grad) {
    cudaError_t e = opt_d.ensure(render_bwd_kernel<F, true>, smem);
    if (e != cudaSuccess) return e;
    launch_pdl(render_bwd_kernel<F, true>, dim3(gx * gy), dim3(RB_THREADS), smem, st, ranges, point_list, rec, W, H, gx, bg, final_T, n_contrib, tile_max, dpix, dpix_F, dpix_D, grad_rec, tile_order);
  } else {
    cudaError_t e = opt_n.ensure(render_bwd_kernel<F, false>, smem);
    if (e != cudaSuccess) return e;
    launch_pdl(render_bwd_kernel<F, false>, dim3(gx * gy), dim3(RB_THREADS), smem, st, ranges, point_list, rec, W, H, gx, bg, final_T, n_contrib, tile_max, dpix, dpix_F, dpix_D, grad_rec, tile_order);
  }
  return cudaGetLastError();
}

cudaError_t launch_render_bwd(int F, const uint2* ranges, const uint32_t* point_list, const float4* rec, int W, int H,
                              int gx, int gy, const float* bg, const float* final_T, const uint32_t* n_contrib,
                              const uint32_t* tile_max, const float* dpix, const float* dpix_F, const float* dpix_D,
                              float* grad_rec, const uint32_t* tile_order, cudaStream_t st) {
  const bool dg = dpix_D != nullptr;
#define RB_CASE(N)                                                                                                \
  case N:                                                                                                         \
    return launch_rb<N>(dg, ranges, point_list, rec, W, H, gx, gy, bg, final_T, n_contrib, tile_max, dpix, dpix_F, \
                        dpix_D, grad_rec, tile_order, st);
  switch (F) {
    RB_CASE(0) RB_CASE(1) RB_CASE(2) RB_CASE(3) RB_CASE(4) RB_CASE(5) RB_CASE(6)
    default: return cudaErrorInvalidValue;
  }
#undef RB_CASE
}

}  // namespace fdls
