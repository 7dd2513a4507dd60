// common.cuh -- shared device helpers for the fdls kernels (sm_100a only).
//
// Numerical contract: the expressions below keep the operation ORDER of the reference device code
// (cuda_rasterizer/auxiliary.h:41-117 and the header-only vec/mat library it is written against), because
// sort keys, tile ranges and per-pixel contributor counts must come out bit-identical (BASELINE.json).
// The file is compiled WITHOUT --use_fast_math and with the default -fmad=true, like the reference
// (setup.py:33 passes no math flags).
#pragma once
#include <atomic>
#include <cstdint>
#include <cstdlib>
#include <mutex>
#include <cuda_runtime.h>
#include "../../include/fdls_abi.h"

namespace fdls {

constexpr int TILE = FDLS_TILE;
constexpr int TILE_PIX = TILE * TILE;
constexpr int REC_FLOATS = 16;  // per-Gaussian render record (64 B): see FdlsLayout::rec
// record slots
enum { R_X = 0, R_Y = 1, R_CA = 2, R_CB = 3, R_CC = 4, R_OP = 5, R_DEPTH = 6, R_RGB = 7, R_LANG = 10 };
// per-Gaussian screen-space gradient record (64 B), accumulated by render_bwd, consumed by preprocess_bwd.
// Slot order follows what the lanes of a quad hold after render_bwd's tensor-core reduction (render_bwd.cu): the first
// 32-byte sector is the "channel" block (sum of weight x dL_dpixel per colour / depth / language channel), the second
// the geometry block + the two language channels that do not fit the first.
enum { G_RGB = 0, G_DEPTH = 3, G_LANG = 4, G_MX = 8, G_MY = 9, G_CA = 10, G_CB = 11, G_CC = 12, G_OP = 13, G_LANG_HI = 14 };
__host__ __device__ constexpr int g_lang_slot(int k) { return k < 4 ? G_LANG + k : G_LANG_HI + (k - 4); }
__host__ __device__ constexpr uint32_t g_used_slots(int F, bool depth) {
  uint32_t m = 0x7u | (0x3fu << G_MX) | (depth ? 1u << G_DEPTH : 0u);
  for (int k = 0; k < F; ++k) m |= 1u << g_lang_slot(k);
  return m;
}

struct V3 {
  float x, y, z;
};
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
__device__ __forceinline__ V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
__device__ __forceinline__ V3& operator+=(V3& a, V3 b) {
  a.x += b.x; a.y += b.y; a.z += b.z;
  return a;
}
__device__ __forceinline__ float dot3(V3 a, V3 b) {
  float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z;
  return tx + ty + tz;
}

// Column-major 3x3 (m[col][row]), product in the same term order as the reference's math library.
struct M3 {
  float m[3][3];
};
__device__ __forceinline__ M3 mul(const M3& a, const M3& b) {
  M3 r;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int k = 0; k < 3; ++k) r.m[c][k] = a.m[0][k] * b.m[c][0] + a.m[1][k] * b.m[c][1] + a.m[2][k] * b.m[c][2];
  return r;
}
__device__ __forceinline__ M3 transpose(const M3& a) {
  M3 r;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int k = 0; k < 3; ++k) r.m[c][k] = a.m[k][c];
  return r;
}

// auxiliary.h:22-39
__device__ constexpr float SH_C0 = 0.28209479177387814f;
__device__ constexpr float SH_C1 = 0.4886025119029199f;
__device__ constexpr float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                       -1.0925484305920792f, 0.5462742152960396f};
__device__ constexpr float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                       0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                       -0.5900435899266435f};

// auxiliary.h:41-44 -- evaluated in double, as the reference's untyped literals force.
__device__ __forceinline__ float ndc2pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

// auxiliary.h:46-56 (radius is an int there; grid = tile grid)
__device__ __forceinline__ void get_rect(float px, float py, int max_radius, int gx, int gy, uint2& rmin, uint2& rmax) {
  rmin.x = min(gx, max((int)0, (int)((px - max_radius) / TILE)));
  rmin.y = min(gy, max((int)0, (int)((py - max_radius) / TILE)));
  rmax.x = min(gx, max((int)0, (int)((px + max_radius + TILE - 1) / TILE)));
  rmax.y = min(gy, max((int)0, (int)((py + max_radius + TILE - 1) / TILE)));
}

// auxiliary.h:58-77
__device__ __forceinline__ float3 xform4x3(const float3& p, const float* m) {
  return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
          m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
}
__device__ __forceinline__ float4 xform4x4(const float3& p, const float* m) {
  return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
          m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14], m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]};
}
__device__ __forceinline__ float3 xform_vec4x3_T(const float3& p, const float* m) {
  return {m[0] * p.x + m[1] * p.y + m[2] * p.z, m[4] * p.x + m[5] * p.y + m[6] * p.z,
          m[8] * p.x + m[9] * p.y + m[10] * p.z};
}
// auxiliary.h:107-117
__device__ __forceinline__ float3 dnormvdv(float3 v, float3 dv) {
  float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
  float invsum32 = 1.0f / sqrt(sum2 * sum2 * sum2);
  float3 o;
  o.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
  o.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
  o.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
  return o;
}

// Per-view constants.  The matrices stay device pointers (they are CUDA tensors in the reference API,
// __init__.py:167-180); kernels stage them in shared memory once per block.
struct ViewParams {
  const float* view;    // (16)
  const float* proj;    // (16)
  const float* campos;  // (3)
  const float* bg;      // (3)
  float tanfovx, tanfovy, focal_x, focal_y, scale_modifier;
  int W, H, gx, gy;
};

// Conservative per-tile visibility summary of one Gaussian, computed once by the thread that stages it:
//   .x = thr : alpha = opacity*exp(power) is certainly < 1/255 whenever power < thr
//   .y = bits: bit w set iff the 16x2-pixel strip of warp w (rows 2w, 2w+1 of the tile) can contain a pixel with
//              power >= thr.  Uses max_dx power(dx,dy) = -0.5*dy^2*det/A (and the same in x for the whole tile).
// Only ever used to SKIP work whose result the exact reference tests would discard anyway; the slack terms (1.001
// relative, +0.05 px, 0.01 in the log domain) are orders of magnitude above fp32 rounding of the exact expression.
__device__ __forceinline__ float2 strip_summary(float gx_, float gy_, float A, float B, float Cc, float op, float tile_x0,
                                                float tile_y0) {
  const float thr = -__logf(255.0f * op) - 0.01f;
  uint32_t bits = 0xffu;
  const float det = A * Cc - B * B;
  if (thr >= 0.0f) {
    bits = 0;  // opacity below 1/255: can never pass the alpha test (power <= 0 for every evaluated pair)
  } else if (det > 0.0f && A > 0.0f && Cc > 0.0f) {
    const float k = -2.0f * thr / det * 1.001f;
    const float ymax = sqrtf(k * A) + 0.05f, xmax = sqrtf(k * Cc) + 0.05f;
    if (gx_ + xmax < tile_x0 || gx_ - xmax > tile_x0 + (TILE - 1)) {
      bits = 0;
    } else {
      const float lo = gy_ - ymax - tile_y0, hi = gy_ + ymax - tile_y0;  // admissible rows, tile-relative
      bits = 0;
#pragma unroll
      for (int w = 0; w < TILE / 2; ++w)
        if (hi >= (float)(2 * w) && lo <= (float)(2 * w + 1)) bits |= 1u << w;
    }
  }  // otherwise (degenerate / NaN conic): keep every strip
  return make_float2(thr, __uint_as_float(bits));
}

// Exact per-block visibility of one Gaussian for a tile cut into eight 8x4-pixel blocks (render_bwd.cu: warp w owns
// block bx = w & 1, by = w >> 1): bit w set iff the block's rectangle of pixel centres can contain a point with
// power >= thr, i.e. iff the minimum of the quadratic form q(d) = A dx^2 + 2 B dx dy + C dy^2 (power = -q/2) over the
// rectangle is <= -2 thr.  q is convex with its minimum at the Gaussian's centre, so over a rectangle that does not
// contain the centre the minimum lies on an edge facing it: on the near vertical edge at dy = clamp(-B dx / C) and on
// the near horizontal edge at dx = clamp(-B dy / A).  (The bounding box of the level set, used by strip_summary, keeps
// ~40 % more (warp, Gaussian) pairs for rotated, elongated Gaussians.)  Same slack terms as strip_summary; only ever
// used to SKIP work the exact reference tests would discard.
// `bx` selects the block column evaluated by this call (two lanes share a Gaussian and OR their results).
__device__ __forceinline__ uint32_t block_summary(float gx_, float gy_, float A, float B, float Cc, float op, float tile_x0,
                                                  float tile_y0, int bx) {
  const float thr = -__logf(255.0f * op) - 0.01f;
  if (thr >= 0.0f) return 0u;  // opacity below 1/255
  const float det = A * Cc - B * B;
  if (!(det > 0.0f && A > 0.0f && Cc > 0.0f)) return 0x55u << bx;  // degenerate / NaN conic: keep every block
  const float K = -2.0f * thr * 1.001f;
  const float gxr = gx_ - tile_x0, gyr = gy_ - tile_y0;
  const float nB_A = -B / A, nB_C = -B / Cc;
  uint32_t bits = 0;
  {
    const float a0 = gxr - (float)(8 * bx + 7) - 0.05f, a1 = gxr - (float)(8 * bx) + 0.05f;  // d = gauss - pixel
    const float dxc = fminf(fmaxf(0.f, a0), a1);
#pragma unroll
    for (int by = 0; by < 4; ++by) {
      const float b0 = gyr - (float)(4 * by + 3) - 0.05f, b1 = gyr - (float)(4 * by) + 0.05f;
      const float dyc = fminf(fmaxf(0.f, b0), b1);
      const float dy1 = fminf(fmaxf(nB_C * dxc, b0), b1);  // best point of the edge dx = dxc
      const float q1 = A * dxc * dxc + 2.f * B * dxc * dy1 + Cc * dy1 * dy1;
      const float dx2 = fminf(fmaxf(nB_A * dyc, a0), a1);  // best point of the edge dy = dyc
      const float q2 = A * dx2 * dx2 + 2.f * B * dx2 * dyc + Cc * dyc * dyc;
      // when the centre's column (row) range contains 0 the corresponding edge is the line through the centre and its
      // candidate is still a point of the rectangle, so min(q1, q2) is attained inside the rectangle in every case
      if (fminf(q1, q2) <= K) bits |= 1u << (2 * by + bx);
    }
  }
  return bits;
}

void count_launch(int n = 1);  // api.cu: feeds fdls_kernel_launches()

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE attribute of a kernel.  One of these per kernel
// remembers the size opted in on each device, so a process that drives several GPUs (one host thread per GPU, or
// one thread switching devices) opts in on every one of them.  Thread-safe: lock-free fast path, mutex on the
// (first-launch) slow path so that two racing threads cannot leave a smaller size behind than the one recorded.
struct SmemOptIn {
  static constexpr int MAX_DEV = 64;
  std::atomic<size_t> bytes[MAX_DEV];
  std::mutex mu;
  SmemOptIn() {
    for (auto& b : bytes) b.store(0, std::memory_order_relaxed);
  }
  template <class K>
  cudaError_t ensure(K* kernel, size_t want) {
    if (want <= 48 * 1024) return cudaSuccess;  // within the default limit on every device
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    const bool tracked = dev >= 0 && dev < MAX_DEV;
    if (tracked && bytes[dev].load(std::memory_order_acquire) >= want) return cudaSuccess;
    std::lock_guard<std::mutex> lock(mu);
    if (tracked && bytes[dev].load(std::memory_order_relaxed) >= want) return cudaSuccess;
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)want);
    if (e == cudaSuccess && tracked) bytes[dev].store(want, std::memory_order_release);
    return e;
  }
};

// ---- Programmatic dependent launch (sm_90+, PTX griddepcontrol) -------------------------------------------------------
// A frame is a chain of 11 short dependent kernels (preprocess -> 4 sort passes -> 2 scans -> 3 partition kernels ->
// compositor), several of them latency-bound; between two plain launches the GPU drains the first grid completely
// before it even starts setting up the second one (~2-3 us per boundary).  Every kernel of the chain is launched with
// programmaticStreamSerializationAllowed and begins with pdl_grid_sync() = griddepcontrol.wait: the launch of kernel
// k+1 is processed while kernel k still runs, and its CTAs start the moment k has completed and flushed.
// Transitivity holds because every kernel waits before it does anything else (a kernel that skipped the wait could
// finish, and release its successor, before an older grid is done).  Launched without the attribute it is a no-op.
// Measured on cfg2 (scripts/ab_pdl.sh, profiles/r02_ab_pdl.txt): one-stream sweep 3287 -> 3601 frames/s, cfg3
// training step 369 -> 385 iters/s.  Also issuing griddepcontrol.launch_dependents at the top of every kernel
// (FDLS_PDL_EARLY=1: successor CTAs become resident and wait on the SM while the predecessor drains) LOST:
// 2882 frames/s -- the resident waiters of up to four younger kernels take slots and shared memory from the one
// that is running.
#ifndef FDLS_PDL_EARLY
#define FDLS_PDL_EARLY 0
#endif
__device__ __forceinline__ void pdl_grid_sync() {
#if FDLS_PDL_EARLY
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
template <class... KArgs, class... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                     Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  static const int pdl = [] { const char* e = getenv("FDLS_PDL"); return e ? atoi(e) : 1; }();  // A/B switch
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

static inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }
static inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

}  // namespace fdls
