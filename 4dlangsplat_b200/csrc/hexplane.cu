// hexplane.cu -- the HexPlane feature field (first step), forward and backward as one kernel each.
// Replaces HexPlaneField.get_density / interpolate_ms_features / grid_sample_wrapper (scene/hexplane.py:22-47,65-106,
// 145-165): per level, product over the six coordinate planes of a bilinear, border-padded, align_corners=True
// grid sample; levels concatenated.  The bilinear arithmetic follows torch's grid_sampler_2d (unnormalise, clip,
// nw/ne/sw/se weights, accumulation order) so the forward agrees with the reference to the last bit or two.
//
// Work split: C/4 consecutive lanes per point, each lane owns one float4 of channels.  Planes are read from
// channel-last (H,W,C) copies, so a corner is one 16-byte load per lane and 4*C contiguous bytes per point; the
// plane set (~30 MB at the HyperNeRF configuration) lives in the 126 MB L2, the only HBM streams are the points in
// and the features out (12 + 4 + 4*L*C bytes per point).
#include "common.cuh"
#include "../../include/fdls_abi.h"

namespace fdls {

constexpr int HX_THREADS = 256;
// itertools.combinations(range(4), 2): xy, xz, xt, yz, yt, zt (hexplane.py:76-78); first index -> width, second -> height
__host__ __device__ constexpr int hx_a0(int ci) { return ci < 3 ? 0 : (ci < 5 ? 1 : 2); }
__host__ __device__ constexpr int hx_a1(int ci) { return ci < 3 ? ci + 1 : (ci < 5 ? ci - 1 : 3); }

struct HxAxis {
  float ix;    // unnormalised, border-clipped coordinate
  float mult;  // d ix / d coord: (size-1)/2, or 0 where the clip is active
  int i0;      // floor(ix)
};

// grid_sampler_unnormalize(align_corners=true) + clip_coordinates(_set_grad) of torch's GridSampler
__device__ __forceinline__ HxAxis hx_axis(float coord, int size) {
  HxAxis r;
  r.ix = ((coord + 1.f) / 2) * (size - 1);
  r.mult = (size - 1) / 2.f;
  if (r.ix <= 0.f) {
    r.ix = 0.f;
    r.mult = 0.f;
  } else {
    const float mx = (float)(size - 1);
    if (r.ix >= mx) {
      r.ix = mx;
      r.mult = 0.f;
    }
  }
  r.i0 = (int)floorf(r.ix);
  return r;
}

__device__ __forceinline__ float4 hx_ld(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 hx_mul(float4 a, float4 b) { return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
__device__ __forceinline__ void hx_fma(float4& acc, float4 v, float w) {
  acc.x += v.x * w;
  acc.y += v.y * w;
  acc.z += v.z * w;
  acc.w += v.w * w;
}

struct HxCorner {
  float nw, ne, sw, se;  // weights
  bool x1, y1;           // east column / south row inside the plane
  int64_t o00;           // element offset of the nw corner (channel 0)
  int64_t dx, dy;        // element strides to east / south; 0 where that neighbour is outside the plane
};

// torch skips a corner that lies outside the plane.  With border clipping that only happens for the east column when
// ix == W-1 exactly (and the south row when iy == H-1), where that corner's weight is exactly 0: reading the clamped
// (= same) texel and adding value*0 gives the same sum for finite planes, and keeps the four loads unconditional so
// all 24 of a level are in flight together.
__device__ __forceinline__ HxCorner hx_corner(const HxAxis& u, const HxAxis& v, int Wd, int Hd, int C) {
  HxCorner c;
  const float x0 = (float)u.i0, y0 = (float)v.i0, x1 = x0 + 1.f, y1 = y0 + 1.f;
  c.nw = (x1 - u.ix) * (y1 - v.ix);
  c.ne = (u.ix - x0) * (y1 - v.ix);
  c.sw = (x1 - u.ix) * (v.ix - y0);
  c.se = (u.ix - x0) * (v.ix - y0);
  c.x1 = u.i0 + 1 < Wd;
  c.y1 = v.i0 + 1 < Hd;
  c.o00 = ((int64_t)v.i0 * Wd + u.i0) * C;
  c.dx = c.x1 ? C : 0;
  c.dy = c.y1 ? (int64_t)Wd * C : 0;
  return c;
}

__device__ __forceinline__ float4 hx_bilinear(const float* plane, const HxCorner& c) {
  const float* b = plane + c.o00;
  const float4 vnw = hx_ld(b), vne = hx_ld(b + c.dx), vsw = hx_ld(b + c.dy), vse = hx_ld(b + c.dy + c.dx);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  hx_fma(acc, vnw, c.nw);
  hx_fma(acc, vne, c.ne);
  hx_fma(acc, vsw, c.sw);
  hx_fma(acc, vse, c.se);
  return acc;
}

// normalize_aabb (hexplane.py:20-21) as the three separate roundings torch performs (no FMA contraction)
__device__ __forceinline__ void hx_coords(const FdlsHexArgs& a, int p, float (&c)[4], float (&scale)[3]) {
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float lo = __ldg(a.aabb + k), hi = __ldg(a.aabb + 3 + k);
    scale[k] = 2.0f / __fsub_rn(hi, lo);
    c[k] = __fsub_rn(__fmul_rn(__fsub_rn(__ldg(a.pts + 3 * (size_t)p + k), lo), scale[k]), 1.0f);
  }
  c[3] = __ldg(a.time + p);
}

template <int G>
__global__ void __launch_bounds__(HX_THREADS) hex_fwd_kernel(const __grid_constant__ FdlsHexArgs a) {
  const int64_t t = (int64_t)blockIdx.x * HX_THREADS + threadIdx.x;
  const int p = (int)(t / G), g = (int)(t % G);
  if (p >= a.P) return;
  const int C = 4 * G;
  float c[4], scale[3];
  hx_coords(a, p, c, scale);
  for (int l = 0; l < a.L; ++l) {
    HxAxis ax[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) ax[k] = hx_axis(c[k], a.reso[l][k]);
    float4 f = make_float4(1.f, 1.f, 1.f, 1.f);
#pragma unroll
    for (int ci = 0; ci < 6; ++ci) {
      const int a0 = hx_a0(ci), a1 = hx_a1(ci);
      const HxCorner cr = hx_corner(ax[a0], ax[a1], a.reso[l][a0], a.reso[l][a1], C);
      f = hx_mul(f, hx_bilinear(a.planes[l][ci] + 4 * g, cr));
    }
    *reinterpret_cast<float4*>(a.out + ((size_t)p * a.L + l) * C + 4 * g) = f;
  }
}

__device__ __forceinline__ float hx_dot(float4 a, float4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
__device__ __forceinline__ float4 hx_scale(float4 a, float s) { return make_float4(a.x * s, a.y * s, a.z * s, a.w * s); }

template <int G>
__global__ void __launch_bounds__(HX_THREADS) hex_bwd_kernel(const __grid_constant__ FdlsHexArgs a) {
  const int64_t t = (int64_t)blockIdx.x * HX_THREADS + threadIdx.x;
  const bool valid = t / G < a.P;
  const int p = valid ? (int)(t / G) : a.P - 1, g = (int)(t % G);
  const int C = 4 * G;
  float c[4], scale[3];
  hx_coords(a, p, c, scale);
  float dcoord[4] = {0.f, 0.f, 0.f, 0.f};
  const bool want_coord = a.dL_dpts != nullptr || a.dL_dtime != nullptr;
  for (int l = 0; l < a.L; ++l) {
    HxAxis ax[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) ax[k] = hx_axis(c[k], a.reso[l][k]);
    // pass 1: the six interpolated vectors
    float4 f[6];
#pragma unroll
    for (int ci = 0; ci < 6; ++ci) {
      const int a0 = hx_a0(ci), a1 = hx_a1(ci);
      const HxCorner cr = hx_corner(ax[a0], ax[a1], a.reso[l][a0], a.reso[l][a1], C);
      f[ci] = hx_bilinear(a.planes[l][ci] + 4 * g, cr);
    }
    // pass 2: d f_i = dL_dout * prod_{j != i} f_j (recomputed per plane: 4 multiplies are cheaper than holding the
    // prefix products in registers), then the plane atomics and the coordinate gradients
    const float4 go = valid ? *reinterpret_cast<const float4*>(a.dL_dout + ((size_t)p * a.L + l) * C + 4 * g)
                            : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int ci = 5; ci >= 0; --ci) {
      const int a0 = hx_a0(ci), a1 = hx_a1(ci);
      float4 gf = go;
#pragma unroll
      for (int cj = 5; cj >= 0; --cj)
        if (cj != ci) gf = hx_mul(gf, f[cj]);
      const HxCorner cr = hx_corner(ax[a0], ax[a1], a.reso[l][a0], a.reso[l][a1], C);
      float* gp = a.dL_dplanes[l][ci];
      if (gp != nullptr && valid) {
        float4* b = reinterpret_cast<float4*>(gp + cr.o00 + 4 * g);
        atomicAdd(b, hx_scale(gf, cr.nw));
        if (cr.x1) atomicAdd(b + cr.dx / 4, hx_scale(gf, cr.ne));
        if (cr.y1) atomicAdd(b + cr.dy / 4, hx_scale(gf, cr.sw));
        if (cr.x1 && cr.y1) atomicAdd(b + (cr.dy + cr.dx) / 4, hx_scale(gf, cr.se));  // dx, dy != 0 here
      }
      if (want_coord) {
        const float* b = a.planes[l][ci] + cr.o00 + 4 * g;
        const HxAxis &u = ax[a0], &v = ax[a1];
        const float x0 = (float)u.i0, y0 = (float)v.i0, x1 = x0 + 1.f, y1 = y0 + 1.f;
        float gix = 0.f, giy = 0.f;
        const float dnw = hx_dot(hx_ld(b), gf), dne = hx_dot(hx_ld(b + cr.dx), gf);
        const float dsw = hx_dot(hx_ld(b + cr.dy), gf), dse = hx_dot(hx_ld(b + cr.dy + cr.dx), gf);
        gix -= dnw * (y1 - v.ix);
        giy -= dnw * (x1 - u.ix);
        if (cr.x1) {
          gix += dne * (y1 - v.ix);
          giy -= dne * (u.ix - x0);
        }
        if (cr.y1) {
          gix -= dsw * (v.ix - y0);
          giy += dsw * (x1 - u.ix);
        }
        if (cr.x1 && cr.y1) {
          gix += dse * (v.ix - y0);
          giy += dse * (u.ix - x0);
        }
        dcoord[a0] += u.mult * gix;
        dcoord[a1] += v.mult * giy;
      }
    }
  }
  if (want_coord) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
#pragma unroll
      for (int o = G / 2; o > 0; o >>= 1) dcoord[k] += __shfl_xor_sync(0xffffffffu, dcoord[k], o);
    }
    if (valid && g == 0) {
      if (a.dL_dpts != nullptr) {
#pragma unroll
        for (int k = 0; k < 3; ++k) a.dL_dpts[3 * (size_t)p + k] = dcoord[k] * scale[k];
      }
      if (a.dL_dtime != nullptr) a.dL_dtime[p] = dcoord[3];
    }
  }
}

// (rows = C) x (cols = HW) row-major  <->  HW x C row-major for up to 24 planes in ONE launch (a level set has 18;
// one 5 us launch per plane would cost more than the backward's atomics).  256 columns per block through a padded
// shared tile; blk0[] is the running block count, so a block finds its plane with a short scan.
struct HxBatch {
  int n, C;
  int64_t hw[FDLS_HEX_MAX_LEVELS * FDLS_HEX_PLANES];
  const float* in[FDLS_HEX_MAX_LEVELS * FDLS_HEX_PLANES];
  float* out[FDLS_HEX_MAX_LEVELS * FDLS_HEX_PLANES];
  int blk0[FDLS_HEX_MAX_LEVELS * FDLS_HEX_PLANES + 1];
};

template <bool PACK>
__global__ void __launch_bounds__(HX_THREADS) hex_transpose_kernel(const __grid_constant__ HxBatch b) {
  extern __shared__ float s_t[];  // [256][C+1]
  int k = 0;
  while (k + 1 < b.n && (int)blockIdx.x >= b.blk0[k + 1]) ++k;
  const int C = b.C;
  const int64_t HW = b.hw[k];
  const float* __restrict__ in = b.in[k];
  float* __restrict__ out = b.out[k];
  const int64_t hw0 = (int64_t)(blockIdx.x - b.blk0[k]) * HX_THREADS;
  const int n = (int)min((int64_t)HX_THREADS, HW - hw0);
  const int tid = threadIdx.x;
  if (PACK) {
    if (tid < n)
      for (int ch = 0; ch < C; ++ch) s_t[tid * (C + 1) + ch] = in[(size_t)ch * HW + hw0 + tid];
    __syncthreads();
    for (int i = tid; i < n * C; i += HX_THREADS) out[(size_t)hw0 * C + i] = s_t[(i / C) * (C + 1) + (i % C)];
  } else {
    for (int i = tid; i < n * C; i += HX_THREADS) s_t[(i / C) * (C + 1) + (i % C)] = in[(size_t)hw0 * C + i];
    __syncthreads();
    if (tid < n)
      for (int ch = 0; ch < C; ++ch) out[(size_t)ch * HW + hw0 + tid] = s_t[tid * (C + 1) + ch];
  }
}

static int hx_check(const FdlsHexArgs* a, bool bwd) {
  if (!a || a->P < 0 || a->L < 1 || a->L > FDLS_HEX_MAX_LEVELS) return FDLS_EINVAL;
  if (a->C != 4 && a->C != 8 && a->C != 16 && a->C != 32) return FDLS_EUNSUPPORTED;
  if (a->P == 0) return FDLS_OK;
  if (!a->pts || !a->time || !a->aabb) return FDLS_EINVAL;
  if (!bwd && !a->out) return FDLS_EINVAL;
  if (bwd && !a->dL_dout) return FDLS_EINVAL;
  for (int l = 0; l < a->L; ++l) {
    for (int k = 0; k < 4; ++k)
      if (a->reso[l][k] < 1) return FDLS_EINVAL;
    for (int ci = 0; ci < FDLS_HEX_PLANES; ++ci)
      if (!a->planes[l][ci]) return FDLS_EINVAL;
  }
  return FDLS_OK;
}

template <bool BWD>
static int hx_launch(const FdlsHexArgs* a, cudaStream_t st) {
  const int G = a->C / 4;
  const int blocks = cdiv((int64_t)a->P * G, HX_THREADS);
  count_launch();
#define HX_CASE(N)                                                \
  case N:                                                         \
    if (BWD)                                                      \
      hex_bwd_kernel<N><<<blocks, HX_THREADS, 0, st>>>(*a);       \
    else                                                          \
      hex_fwd_kernel<N><<<blocks, HX_THREADS, 0, st>>>(*a);       \
    break;
  switch (G) { HX_CASE(1) HX_CASE(2) HX_CASE(4) HX_CASE(8) }
#undef HX_CASE
  return cudaGetLastError() == cudaSuccess ? FDLS_OK : FDLS_ECUDA;
}

int hexplane_forward(const FdlsHexArgs* a, cudaStream_t st) {
  const int rc = hx_check(a, false);
  if (rc != FDLS_OK || a->P == 0) return rc;
  return hx_launch<false>(a, st);
}

int hexplane_backward(const FdlsHexArgs* a, cudaStream_t st) {
  const int rc = hx_check(a, true);
  if (rc != FDLS_OK || a->P == 0) return rc;
  return hx_launch<true>(a, st);
}

int hexplane_transpose(bool pack, int n, int C, const int64_t* HW, const float* const* in, float* const* out,
                       cudaStream_t st) {
  if (C < 1 || C > 64 || n < 0 || n > FDLS_HEX_MAX_LEVELS * FDLS_HEX_PLANES) return FDLS_EINVAL;
  if (n > 0 && (!HW || !in || !out)) return FDLS_EINVAL;
  HxBatch b;
  b.n = 0;
  b.C = C;
  b.blk0[0] = 0;
  for (int i = 0; i < n; ++i) {
    if (HW[i] < 0) return FDLS_EINVAL;
    if (HW[i] == 0) continue;
    if (!in[i] || !out[i]) return FDLS_EINVAL;
    b.hw[b.n] = HW[i];
    b.in[b.n] = in[i];
    b.out[b.n] = out[i];
    b.blk0[b.n + 1] = b.blk0[b.n] + cdiv(HW[i], HX_THREADS);
    ++b.n;
  }
  if (b.n == 0) return FDLS_OK;
  const size_t smem = (size_t)HX_THREADS * (C + 1) * sizeof(float);
  count_launch();
  if (pack)
    hex_transpose_kernel<true><<<b.blk0[b.n], HX_THREADS, smem, st>>>(b);
  else
    hex_transpose_kernel<false><<<b.blk0[b.n], HX_THREADS, smem, st>>>(b);
  return cudaGetLastError() == cudaSuccess ? FDLS_OK : FDLS_ECUDA;
}

}  // namespace fdls
