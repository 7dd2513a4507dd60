// binning_bucket.cu -- tile-bucketed binning: the default replacement of duplicateWithKeys + cub SortPairs +
// identifyTileRanges (rasterizer_impl.cu:70-138, :293-318).
//
// The reference builds R (tile<<32|depth) keys and runs a 6-pass LSD radix sort over all R pairs.  But the high key
// field is a small integer (2k-8k tiles) and the final order is simply "per tile, by (depth bits, Gaussian index)":
//   K1 tile_count    one RED per (Gaussian, tile) overlap into T counters
//   K2 tile_scan     one CTA: exclusive scan over the T counters -> tile starts == the reference's `ranges`, and R
//   K3 tile_scatter  each overlap takes a slot in its tile's segment (atomic cursor) and stores (depth<<32 | index)
//   K4 tile_sort     one CTA per tile sorts its segment IN SHARED MEMORY (bitonic network on the unique 64-bit
//                    (depth, index) keys; segments longer than 4096 are sorted in 4096-element runs and merged with
//                    merge-path passes), then writes point_list and the reference-format keys.
// Because (depth bits, index) is a total order, the result is IDENTICAL to the reference's stable radix sort for
// any scatter order (ties in depth keep ascending Gaussian index exactly as stability does there), so sorted keys /
// values / ranges stay bit-exact, while the HBM traffic drops from ~150 B to ~40 B per instance and no pass over R is
// serialised behind a global digit histogram.  The onesweep sort (radix_sort.cu) remains selectable (binning_algo=1).
#include "common.cuh"

namespace fdls {

constexpr int TS_MAX = 4096;     // segment length sorted in one go in shared memory (32 KB of u64)
constexpr int TS_THREADS = 256;

// Visit every (tile) of every visible Gaussian; warp-cooperative for Gaussians covering many tiles.
template <bool SCATTER>
__global__ void __launch_bounds__(256)
tile_visit_kernel(int P, const float4* __restrict__ rec, const int* __restrict__ radii, int gx, int gy,
                  uint32_t* __restrict__ counters, const uint32_t* __restrict__ tile_start, uint64_t* __restrict__ seg,
                  long long capacity) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  const int radius = idx < P ? radii[idx] : 0;
  uint32_t depth_bits = 0, count = 0;
  uint2 rmin = {0, 0}, rmax = {0, 0};
  if (radius > 0) {
    const float4 r0 = rec[(size_t)idx * 4];
    if (SCATTER) depth_bits = __float_as_uint(rec[(size_t)idx * 4 + 1].z);
    get_rect(r0.x, r0.y, radius, gx, gy, rmin, rmax);
    count = (rmax.x - rmin.x) * (rmax.y - rmin.y);
  }
  auto visit = [&](uint32_t tile, uint32_t d_bits, uint32_t id) {
    if (SCATTER) {
      const uint32_t slot = tile_start[tile] + atomicAdd(&counters[tile], 1u);
      if ((long long)slot < capacity) seg[slot] = ((uint64_t)d_bits << 32) | id;
    } else {
      atomicAdd(&counters[tile], 1u);  // result unused -> RED
    }
  };
  constexpr uint32_t SMALL = 4;
  if (count > 0 && count <= SMALL) {
    for (uint32_t y = rmin.y; y < rmax.y; y++)
      for (uint32_t x = rmin.x; x < rmax.x; x++) visit(y * gx + x, depth_bits, (uint32_t)idx);
  }
  uint32_t big = __ballot_sync(0xffffffffu, count > SMALL);
  while (big) {
    const int src = __ffs(big) - 1;
    big &= big - 1;
    const uint32_t b_cnt = __shfl_sync(0xffffffffu, count, src);
    const uint32_t b_x0 = __shfl_sync(0xffffffffu, rmin.x, src);
    const uint32_t b_y0 = __shfl_sync(0xffffffffu, rmin.y, src);
    const uint32_t b_w = __shfl_sync(0xffffffffu, rmax.x - rmin.x, src);
    const uint32_t b_depth = __shfl_sync(0xffffffffu, depth_bits, src);
    const uint32_t b_idx = __shfl_sync(0xffffffffu, (uint32_t)idx, src);
    for (uint32_t i = lane; i < b_cnt; i += 32) visit((b_y0 + i / b_w) * gx + b_x0 + i % b_w, b_depth, b_idx);
  }
}

// One CTA.  tile_start[t] = exclusive prefix, tile_start[T] = R; ranges in the reference's convention (empty tiles
// stay (0,0): rasterizer_impl.cu:314 memset + :116-138); counters are reset to serve as K3's cursors.
__global__ void __launch_bounds__(1024)
tile_scan_kernel(int T, uint32_t* __restrict__ counters, uint32_t* __restrict__ tile_start, uint2* __restrict__ ranges,
                 uint32_t* __restrict__ num_rendered, long long capacity) {
  __shared__ uint32_t s_warp[33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t carry = 0;
  for (int base = 0; base < T; base += 1024) {
    const int t = base + threadIdx.x;
    const uint32_t c = t < T ? counters[t] : 0;
    uint32_t incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      const uint32_t w = s_warp[lane];
      uint32_t wi = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, wi, d);
        if (lane >= d) wi += v;
      }
      s_warp[lane] = wi - w;
      if (lane == 31) s_warp[32] = wi;
    }
    __syncthreads();
    const uint32_t start = carry + s_warp[warp] + incl - c;
    if (t < T) {
      tile_start[t] = start;
      ranges[t] = c ? make_uint2(start, start + c) : make_uint2(0u, 0u);
      counters[t] = 0;
    }
    carry += s_warp[32];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    tile_start[T] = carry;
    num_rendered[0] = carry;
    num_rendered[1] = ((long long)carry > capacity) ? 1u : 0u;
  }
}

__device__ __forceinline__ void bitonic_sort_smem(uint64_t* a, int N) {  // N: power of two, ascending
  for (int k = 2; k <= N; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < N; i += TS_THREADS) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const uint64_t x = a[i], y = a[ixj];
          const bool asc = (i & k) == 0;
          if ((x > y) == asc) {
            a[i] = y;
            a[ixj] = x;
          }
        }
      }
      __syncthreads();
    }
  }
}

// Merge the sorted runs src[a0,a1) and src[b0,b1) into dst starting at a0, cooperatively (merge-path).
// (plain pointers on purpose: both buffers are written earlier in this kernel, so no read-only/non-coherent loads)
__device__ void merge_runs(uint64_t* src, uint64_t* dst, int a0, int a1, int b0, int b1) {
  const int na = a1 - a0, nb = b1 - b0, total = na + nb;
  constexpr int PER = 8;
  for (int o = threadIdx.x * PER; o < total; o += TS_THREADS * PER) {
    // diagonal o: find i (elements taken from A) with A[i-1] <= B[o-i] and B[o-i-1] < A[i]
    int lo = max(0, o - nb), hi = min(o, na);
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (src[a0 + mid] <= src[b0 + o - 1 - mid]) lo = mid + 1;
      else hi = mid;
    }
    int i = lo, j = o - lo;
    const int end = min(o + PER, total);
    for (int k = o; k < end; ++k) {
      const bool takeA = j >= nb || (i < na && src[a0 + i] <= src[b0 + j]);
      dst[a0 + k] = takeA ? src[a0 + i++] : src[b0 + j++];
    }
  }
}

__global__ void __launch_bounds__(TS_THREADS)
tile_sort_kernel(int T, const uint32_t* __restrict__ tile_start, const uint32_t* __restrict__ num_rendered,
                 long long capacity, uint64_t* seg, uint64_t* keys_out, uint32_t* __restrict__ point_list) {
  __shared__ uint64_t s[TS_MAX];
  const uint32_t tile = blockIdx.x;
  const long long lim = min((long long)num_rendered[0], capacity);
  const long long beg = min((long long)tile_start[tile], lim), fin = min((long long)tile_start[tile + 1], lim);
  const int n = (int)(fin - beg);
  if (n <= 0) return;
  uint64_t* my = seg + beg;
  uint64_t* out = keys_out + beg;
  const uint64_t tile_hi = (uint64_t)tile << 32;
  if (n <= TS_MAX) {
    int N = 32;
    while (N < n) N <<= 1;
    for (int i = threadIdx.x; i < N; i += TS_THREADS) s[i] = i < n ? my[i] : ~0ull;
    __syncthreads();
    bitonic_sort_smem(s, N);
    for (int i = threadIdx.x; i < n; i += TS_THREADS) {
      const uint64_t k = s[i];
      point_list[beg + i] = (uint32_t)k;
      out[i] = tile_hi | (k >> 32);  // the reference's key: tile | depth bits
    }
    return;
  }
  // long segment: sort 4096-element runs in shared memory, then merge runs pairwise, ping-ponging seg <-> keys_out
  for (int c0 = 0; c0 < n; c0 += TS_MAX) {
    const int m = min(TS_MAX, n - c0);
    for (int i = threadIdx.x; i < TS_MAX; i += TS_THREADS) s[i] = i < m ? my[c0 + i] : ~0ull;
    __syncthreads();
    bitonic_sort_smem(s, TS_MAX);
    for (int i = threadIdx.x; i < m; i += TS_THREADS) my[c0 + i] = s[i];
    __syncthreads();
  }
  uint64_t* src = my;
  uint64_t* dst = out;
  for (int width = TS_MAX; width < n; width <<= 1) {
    for (int a0 = 0; a0 < n; a0 += 2 * width) {
      const int a1 = min(a0 + width, n), b1 = min(a0 + 2 * width, n);
      if (a1 < b1) {
        merge_runs(src, dst, a0, a1, a1, b1);
      } else {
        for (int i = a0 + threadIdx.x; i < a1; i += TS_THREADS) dst[i] = src[i];
      }
    }
    __threadfence_block();
    __syncthreads();
    uint64_t* t = src; src = dst; dst = t;
  }
  // src holds the sorted (depth<<32|index) keys; emit ids and reference-format keys (in place when src == out)
  for (int i = threadIdx.x; i < n; i += TS_THREADS) {
    const uint64_t k = src[i];
    point_list[beg + i] = (uint32_t)k;
    out[i] = tile_hi | (k >> 32);
  }
}

cudaError_t launch_tile_count(int P, const float4* rec, const int* radii, int gx, int gy, uint32_t* counters,
                              cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(counters, 0, (size_t)gx * gy * sizeof(uint32_t), st);
  if (e != cudaSuccess) return e;
  if (P > 0) {
    tile_visit_kernel<false><<<cdiv(P, 256), 256, 0, st>>>(P, rec, radii, gx, gy, counters, nullptr, nullptr, 0);
    count_launch();
  }
  return cudaGetLastError();
}

cudaError_t launch_tile_scan(int T, uint32_t* counters, uint32_t* tile_start, uint2* ranges, uint32_t* num_rendered,
                             long long capacity, cudaStream_t st) {
  tile_scan_kernel<<<1, 1024, 0, st>>>(T, counters, tile_start, ranges, num_rendered, capacity);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_tile_scatter_sort(int P, const float4* rec, const int* radii, int gx, int gy, uint32_t* counters,
                                     const uint32_t* tile_start, const uint32_t* num_rendered, long long capacity,
                                     uint64_t* seg, uint64_t* keys_out, uint32_t* point_list, void** events,
                                     cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(counters, 0, (size_t)gx * gy * sizeof(uint32_t), st);  // cursors (re-runs too)
  if (e != cudaSuccess) return e;
  if (P > 0) {
    tile_visit_kernel<true><<<cdiv(P, 256), 256, 0, st>>>(P, rec, radii, gx, gy, counters, tile_start, seg, capacity);
    count_launch();
  }
  if (events && events[3]) cudaEventRecord((cudaEvent_t)events[3], st);
  tile_sort_kernel<<<gx * gy, TS_THREADS, 0, st>>>(gx * gy, tile_start, num_rendered, capacity, seg, keys_out,
                                                   point_list);
  count_launch();
  return cudaGetLastError();
}

}  // namespace fdls
