// binning.cu -- tile-count scan, (tile|depth) key duplication and tile-range detection.
//
// Replaces cub::DeviceScan::InclusiveSum (rasterizer_impl.cu:281), the blocking D2H read of num_rendered (:285),
// duplicateWithKeys (:70-111) and identifyTileRanges (:116-138) + cudaMemset (:314).
//
// B200 notes: the instance count R never has to reach the host -- every kernel reads it from device memory and is
// launched over a capacity-sized grid; duplication is warp-cooperative (a Gaussian covering many tiles is emitted by
// all 32 lanes with coalesced 8-byte/4-byte stores instead of one lane looping while 31 idle).
#include "common.cuh"

namespace fdls {

constexpr int SCAN_THREADS = 512;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;  // 4096 elements per block

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, int lane) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
    if (lane >= d) v += t;
  }
  return v;
}

// block-wide exclusive scan of one value per thread; returns the exclusive prefix, total in `total`
template <int THREADS>
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t& total, uint32_t* s_warp) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = warp_incl_scan(v, lane);
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < THREADS / 32 ? s_warp[lane] : 0;
    uint32_t wi = warp_incl_scan(w, lane);
    s_warp[lane] = wi - w;  // exclusive warp offsets; slot THREADS/32 .. unused
    if (lane == 31) s_warp[32] = wi;
  }
  __syncthreads();
  total = s_warp[32];
  uint32_t r = incl - v + s_warp[warp];
  __syncthreads();
  return r;
}

// Block sums; the LAST block to finish (ticket in num_rendered[2], zeroed by the preprocess kernel and reset here)
// also scans them in place and publishes R and the overflow flag, which saves the one-block spine launch.
__global__ void __launch_bounds__(SCAN_THREADS)
scan_reduce_kernel(int P, const uint32_t* __restrict__ in, const uint32_t* __restrict__ order,
                   uint32_t* __restrict__ block_sums, uint32_t* __restrict__ num_rendered, long long capacity,
                   uint32_t* host_mirror) {
  pdl_grid_sync();
  __shared__ uint32_t s_warp[33];
  __shared__ uint32_t s_last;
  const int base = blockIdx.x * SCAN_TILE;
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    int k = base + i * SCAN_THREADS + threadIdx.x;
    if (k < P) s += in[order ? order[k] : k];
  }
  uint32_t total;
  block_excl_scan<SCAN_THREADS>(s, total, s_warp);
  if (threadIdx.x == 0) {
    block_sums[blockIdx.x] = total;
    __threadfence();
    s_last = atomicAdd(&num_rendered[2], 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const int nblocks = gridDim.x;
  uint32_t carry = 0;
  for (int b0 = 0; b0 < nblocks; b0 += SCAN_THREADS) {
    const int k = b0 + threadIdx.x;
    const uint32_t v = k < nblocks ? __ldcg(&block_sums[k]) : 0;
    uint32_t tot;
    const uint32_t ex = block_excl_scan<SCAN_THREADS>(v, tot, s_warp);
    if (k < nblocks) block_sums[k] = carry + ex;
    carry += tot;
  }
  if (threadIdx.x == 0) {
    num_rendered[0] = carry;
    num_rendered[1] = ((long long)carry > capacity) ? 1u : 0u;
    num_rendered[2] = 0;  // ready for the next use of this workspace
    if (host_mirror != nullptr) {  // pinned, device-mapped host memory: no copy-engine hop between the kernels
      volatile uint32_t* h = host_mirror;
      h[0] = carry;
      h[1] = ((long long)carry > capacity) ? 1u : 0u;
      __threadfence_system();
    }
  }
}

// one block: exclusive scan of the block sums in place; publishes R and the overflow flag
__global__ void __launch_bounds__(1024)
scan_spine_kernel(int nblocks, uint32_t* __restrict__ block_sums, uint32_t* __restrict__ num_rendered,
                  long long capacity, uint32_t* host_mirror) {
  pdl_grid_sync();
  __shared__ uint32_t s_warp[33];
  uint32_t carry = 0;
  for (int base = 0; base < nblocks; base += 1024) {
    int k = base + threadIdx.x;
    uint32_t v = k < nblocks ? block_sums[k] : 0;
    uint32_t total;
    uint32_t ex = block_excl_scan<1024>(v, total, s_warp);
    if (k < nblocks) block_sums[k] = carry + ex;
    carry += total;
  }
  if (threadIdx.x == 0) {
    num_rendered[0] = carry;
    num_rendered[1] = ((long long)carry > capacity) ? 1u : 0u;
    if (host_mirror != nullptr) {
      volatile uint32_t* h = host_mirror;
      h[0] = carry;
      h[1] = ((long long)carry > capacity) ? 1u : 0u;
      __threadfence_system();
    }
  }
}

__global__ void __launch_bounds__(SCAN_THREADS)
scan_apply_kernel(int P, const uint32_t* __restrict__ in, const uint32_t* __restrict__ order,
                  const uint32_t* __restrict__ block_sums, uint32_t* __restrict__ out) {
  pdl_grid_sync();
  __shared__ uint32_t s_warp[33];
  // blocked arrangement: thread t owns items [t*ITEMS, t*ITEMS+ITEMS) of the block tile
  const int base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS];
  uint32_t s = 0;
  if (order != nullptr) {  // gather in depth order
    if (base + SCAN_ITEMS <= P) {
      const uint4* p4 = reinterpret_cast<const uint4*>(order + base);
      const uint4 a = p4[0], b = p4[1];
      const uint32_t id[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < SCAN_ITEMS; ++i) v[i] = in[id[i]];
    } else {
#pragma unroll
      for (int i = 0; i < SCAN_ITEMS; ++i) v[i] = (base + i < P) ? in[order[base + i]] : 0;
    }
  } else if (base + SCAN_ITEMS <= P) {
    const uint4* p4 = reinterpret_cast<const uint4*>(in + base);
    uint4 a = p4[0], b = p4[1];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) v[i] = (base + i < P) ? in[base + i] : 0;
  }
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) s += v[i];
  uint32_t total;
  uint32_t ex = block_excl_scan<SCAN_THREADS>(s, total, s_warp) + block_sums[blockIdx.x];
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    ex += v[i];
    v[i] = ex;
  }
  if (base + SCAN_ITEMS <= P) {
    uint4* o4 = reinterpret_cast<uint4*>(out + base);
    o4[0] = make_uint4(v[0], v[1], v[2], v[3]);
    o4[1] = make_uint4(v[4], v[5], v[6], v[7]);
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i)
      if (base + i < P) out[base + i] = v[i];
  }
}

// ---- single-pass scan (default): reduce + look-back + apply in ONE kernel --------------------------------------------
// The two-kernel scan above loads every element twice through the same dependent chain (order[k], then in[order[k]])
// and pays two launches on the frame's critical path: 11 + 7.5 us at cfg2 for 74 blocks that keep the GPU 6 % busy.
// Here a block takes a tile by ticket (so it only ever waits for blocks that already started), publishes its aggregate
// in a 64-bit state word {flag, value}, obtains its exclusive prefix by decoupled look-back over its predecessors' words
// (one warp, 32 predecessors per step), publishes the inclusive value, and writes its part of the scan.  The state words
// and the two tickets start at zero (the depth sort's scratch memset covers them, see api.cu); the block that finishes
// last re-zeroes the tickets.
constexpr uint64_t SCAN_FLAG_AGG = 1ull << 32, SCAN_FLAG_INC = 2ull << 32;

__global__ void __launch_bounds__(SCAN_THREADS)
scan_onepass_kernel(int P, const uint32_t* __restrict__ in, const uint32_t* __restrict__ order,
                    unsigned long long* __restrict__ state, uint32_t* __restrict__ out,
                    uint32_t* __restrict__ num_rendered, long long capacity, uint32_t* host_mirror) {
  pdl_grid_sync();
  __shared__ uint32_t s_warp[33];
  __shared__ uint32_t s_tile, s_prefix;
  if (threadIdx.x == 0) s_tile = atomicAdd(&num_rendered[2], 1u);
  __syncthreads();
  const int tile = (int)s_tile, nb = gridDim.x;
  // blocked arrangement: thread t owns items [t*ITEMS, t*ITEMS+ITEMS) of the tile
  const int base = tile * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS];
  if (base + SCAN_ITEMS <= P) {
    if (order != nullptr) {  // gather in depth order
      const uint4* p4 = reinterpret_cast<const uint4*>(order + base);
      const uint4 a = p4[0], b = p4[1];
      const uint32_t id[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < SCAN_ITEMS; ++i) v[i] = in[id[i]];
    } else {
      const uint4* p4 = reinterpret_cast<const uint4*>(in + base);
      const uint4 a = p4[0], b = p4[1];
      v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) v[i] = (base + i < P) ? in[order ? order[base + i] : base + i] : 0;
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) s += v[i];
  uint32_t total;
  uint32_t ex = block_excl_scan<SCAN_THREADS>(s, total, s_warp);
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    volatile unsigned long long* vs = state;
    if (lane == 0) {
      vs[tile] = (tile == 0 ? SCAN_FLAG_INC : SCAN_FLAG_AGG) | total;
      __threadfence();
    }
    uint32_t prefix = 0;
    for (int j0 = tile - 1; j0 >= 0; j0 -= 32) {
      const int j = j0 - lane;
      unsigned long long w = SCAN_FLAG_INC;  // lanes before tile 0 read as "inclusive 0"
      if (j >= 0) {
        uint32_t polls = 0;
        do {
          w = vs[j];
          if (++polls > (1u << 26)) __trap();  // a predecessor that never publishes must fail loudly, not hang the GPU
        } while ((w >> 32) == 0);
      }
      const uint32_t inc = __ballot_sync(0xffffffffu, (w & SCAN_FLAG_INC) != 0);
      const int stop = __ffs(inc) - 1;  // nearest predecessor that already knows its inclusive value (-1: none)
      uint32_t val = (stop < 0 || lane <= stop) ? (uint32_t)w : 0u;
#pragma unroll
      for (int d = 16; d >= 1; d >>= 1) val += __shfl_xor_sync(0xffffffffu, val, d);
      prefix += val;
      if (stop >= 0) break;
    }
    if (lane == 0) {
      if (tile > 0) {
        vs[tile] = SCAN_FLAG_INC | (unsigned long long)(prefix + total);
        __threadfence();
      }
      s_prefix = prefix;
    }
  }
  __syncthreads();
  ex += s_prefix;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    ex += v[i];
    v[i] = ex;
  }
  if (base + SCAN_ITEMS <= P) {
    uint4* o4 = reinterpret_cast<uint4*>(out + base);
    o4[0] = make_uint4(v[0], v[1], v[2], v[3]);
    o4[1] = make_uint4(v[4], v[5], v[6], v[7]);
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i)
      if (base + i < P) out[base + i] = v[i];
  }
  if (threadIdx.x == 0) {
    if (tile == nb - 1) {  // the grand total: R and the overflow flag
      const uint32_t R = s_prefix + total;
      num_rendered[0] = R;
      num_rendered[1] = ((long long)R > capacity) ? 1u : 0u;
      if (host_mirror != nullptr) {  // pinned, device-mapped host memory: no copy-engine hop between the kernels
        volatile uint32_t* h = host_mirror;
        h[0] = R;
        h[1] = ((long long)R > capacity) ? 1u : 0u;
        __threadfence_system();
      }
      __threadfence();
    }
    // every block counts itself out on the partition's ticket word (zero between its uses); the last one resets both
    if (atomicAdd(&num_rendered[3], 1u) == (uint32_t)nb - 1) {
      num_rendered[2] = 0;
      num_rendered[3] = 0;
    }
  }
}

size_t scan_state_bytes(int P) { return align_up((size_t)cdiv(P < 1 ? 1 : P, SCAN_TILE) * sizeof(unsigned long long)); }

// rasterizer_impl.cu:70-111, warp-cooperative.  Thread `pos` handles the pos-th Gaussian in DEPTH order
// (order[pos]); within a Gaussian the tiles are emitted y-major, x-minor like the reference.  Because instances
// leave this kernel already depth-sorted, the following stable sort only has to look at the tile bits -- and since the
// tile ids are produced right here, the kernel also accumulates the sort's per-pass digit histograms (shared-memory
// counters, warp-aggregated with match.any), which removes the sort's own histogram pass over the keys.
constexpr int DUP_MAX_PASSES = 3;  // tile ids up to 24 bits

__device__ __forceinline__ void dup_count(uint32_t (*s_hist)[256], int passes, bool valid, uint32_t tile, uint32_t lt) {
#pragma unroll
  for (int p = 0; p < DUP_MAX_PASSES; ++p) {
    if (p < passes) {
      const uint32_t d = valid ? ((tile >> (8 * p)) & 255u) : 0xffffffffu;
      const unsigned peers = __match_any_sync(0xffffffffu, d);
      if (valid && (peers & lt) == 0) atomicAdd(&s_hist[p][d], __popc(peers));
    }
  }
}

__global__ void __launch_bounds__(256)
duplicate_keys_kernel(int P, const float4* __restrict__ rec, const uint32_t* __restrict__ order,
                      const uint32_t* __restrict__ offsets, const int* __restrict__ radii,
                      uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, int gx, int gy, long long capacity,
                      uint32_t* __restrict__ global_hist, int passes) {
  pdl_grid_sync();
  __shared__ uint32_t s_hist[DUP_MAX_PASSES][256];
  for (int i = threadIdx.x; i < DUP_MAX_PASSES * 256; i += blockDim.x) (&s_hist[0][0])[i] = 0;
  __syncthreads();
  const int pos = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  uint32_t lt;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(lt));
  const int idx = pos < P ? (int)order[pos] : 0;
  int radius = pos < P ? radii[idx] : 0;
  uint32_t off = 0, depth_bits = 0, count = 0;
  uint2 rmin = {0, 0}, rmax = {0, 0};
  if (radius > 0) {
    off = (pos == 0) ? 0 : offsets[pos - 1];
    const float4 r0 = rec[(size_t)idx * 4];
    depth_bits = __float_as_uint(rec[(size_t)idx * 4 + 1].z);
    get_rect(r0.x, r0.y, radius, gx, gy, rmin, rmax);
    count = (rmax.x - rmin.x) * (rmax.y - rmin.y);
  }
  constexpr uint32_t SMALL = 4;
  {  // Gaussians touching <= SMALL tiles: each lane emits its own (uniform trip count keeps the warp converged)
    const bool mine = count > 0 && count <= SMALL;
    const uint32_t w = rmax.x - rmin.x;
#pragma unroll
    for (uint32_t k = 0; k < SMALL; ++k) {
      const bool valid = mine && k < count;
      uint32_t tile = 0;
      if (valid) {
        const uint32_t y = rmin.y + k / w, x = rmin.x + k % w;
        tile = y * gx + x;
        const uint32_t o = off + k;
        if ((long long)o < capacity) {
          keys[o] = ((uint64_t)tile << 32) | depth_bits;
          vals[o] = idx;
        } else {
          tile = 0xffffffffu;
        }
      }
      dup_count(s_hist, passes, valid && tile != 0xffffffffu, tile, lt);
    }
  }
  uint32_t big = __ballot_sync(0xffffffffu, count > SMALL);
  while (big) {
    const int src = __ffs(big) - 1;
    big &= big - 1;
    const uint32_t b_off = __shfl_sync(0xffffffffu, off, src);
    const uint32_t b_cnt = __shfl_sync(0xffffffffu, count, src);
    const uint32_t b_x0 = __shfl_sync(0xffffffffu, rmin.x, src);
    const uint32_t b_y0 = __shfl_sync(0xffffffffu, rmin.y, src);
    const uint32_t b_w = __shfl_sync(0xffffffffu, rmax.x - rmin.x, src);
    const uint32_t b_depth = __shfl_sync(0xffffffffu, depth_bits, src);
    const uint32_t b_idx = __shfl_sync(0xffffffffu, (uint32_t)idx, src);
    for (uint32_t base = 0; base < b_cnt; base += 32) {
      const uint32_t i = base + lane;
      bool valid = i < b_cnt;
      uint32_t tile = 0;
      if (valid) {
        const uint32_t y = b_y0 + i / b_w, x = b_x0 + i % b_w;
        tile = y * gx + x;
        const uint32_t o = b_off + i;
        if ((long long)o < capacity) {
          keys[o] = ((uint64_t)tile << 32) | b_depth;
          vals[o] = b_idx;
        } else {
          valid = false;
        }
      }
      dup_count(s_hist, passes, valid, tile, lt);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * 256; i += blockDim.x) {
    const uint32_t c = (&s_hist[0][0])[i];
    if (c) atomicAdd(&global_hist[i], c);
  }
}

// rasterizer_impl.cu:116-138 with the element count read on the device
__global__ void __launch_bounds__(256)
tile_ranges_kernel(const uint32_t* __restrict__ num_rendered, long long capacity, const uint64_t* __restrict__ keys,
                   uint2* __restrict__ ranges) {
  pdl_grid_sync();
  const long long L = min((long long)num_rendered[0], capacity);
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= L) return;
  const uint32_t currtile = keys[idx] >> 32;
  if (idx == 0)
    ranges[currtile].x = 0;
  else {
    const uint32_t prevtile = keys[idx - 1] >> 32;
    if (currtile != prevtile) {
      ranges[prevtile].y = (uint32_t)idx;
      ranges[currtile].x = (uint32_t)idx;
    }
  }
  if (idx == L - 1) ranges[currtile].y = (uint32_t)L;
}

size_t scan_scratch_bytes(int P) { return align_up((size_t)(cdiv(P, SCAN_TILE) + 1) * sizeof(uint32_t)); }

cudaError_t launch_scan(int P, const uint32_t* tiles_touched, const uint32_t* order, uint32_t* offsets,
                        uint32_t* block_sums, unsigned long long* state, uint32_t* num_rendered, long long capacity,
                        uint32_t* host_mirror, cudaStream_t st) {
  const int nb = cdiv(P, SCAN_TILE);
  static const bool onepass = [] {
    const char* e = getenv("FDLS_SCAN");
    return !(e && e[0] == '2');  // FDLS_SCAN=2: the two-kernel scan, for A/B
  }();
  if (P > 0 && onepass && state != nullptr) {
    launch_pdl(scan_onepass_kernel, dim3(nb), dim3(SCAN_THREADS), 0, st, P, tiles_touched, order, state, offsets, num_rendered, capacity, host_mirror);
    count_launch();
    return cudaGetLastError();
  }
  if (P > 0) {
    launch_pdl(scan_reduce_kernel, dim3(nb), dim3(SCAN_THREADS), 0, st, P, tiles_touched, order, block_sums, num_rendered, capacity, host_mirror);
    count_launch();
  } else {
    launch_pdl(scan_spine_kernel, dim3(1), dim3(1024), 0, st, nb, block_sums, num_rendered, capacity, host_mirror); count_launch();
  }
  if (P > 0) { launch_pdl(scan_apply_kernel, dim3(nb), dim3(SCAN_THREADS), 0, st, P, tiles_touched, order, block_sums, offsets); count_launch(); }
  return cudaGetLastError();
}

cudaError_t launch_duplicate(int P, const float4* rec, const uint32_t* order, const uint32_t* offsets,
                             const int* radii, uint64_t* keys, uint32_t* vals, int gx, int gy, long long capacity,
                             uint32_t* global_hist, int passes, cudaStream_t st) {
  if (P == 0) return cudaSuccess;
  if (passes > DUP_MAX_PASSES) return cudaErrorInvalidValue;
  launch_pdl(duplicate_keys_kernel, dim3(cdiv(P, 256)), dim3(256), 0, st, P, rec, order, offsets, radii, keys, vals, gx, gy, capacity, global_hist, passes); count_launch();
  return cudaGetLastError();
}

cudaError_t launch_tile_ranges(const uint32_t* num_rendered, long long n_grid, long long capacity, const uint64_t* keys,
                               uint2* ranges, int tiles, cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(ranges, 0, (size_t)tiles * sizeof(uint2), st);
  if (e != cudaSuccess) return e;
  if (n_grid > 0) { launch_pdl(tile_ranges_kernel, dim3(cdiv(n_grid, 256)), dim3(256), 0, st, num_rendered, capacity, keys, ranges); count_launch(); }
  return cudaGetLastError();
}

}  // namespace fdls
