// binning_partition.cu -- default binning: ONE stable partition pass by tile over depth-ordered Gaussians.
// Replaces duplicateWithKeys + cub::DeviceRadixSort::SortPairs + identifyTileRanges
// (rasterizer_impl.cu:70-138, :293-318) without ever materialising unsorted keys.
//
// The reference sorts R (tile<<32|depth) keys with a 6-pass LSD radix sort.  The same order is obtained by
//   (1) sorting the P GAUSSIANS by depth bits once (stable onesweep over P elements, radix_sort.cu; R/P ~ 10), and
//   (2) a stable multi-way partition of the depth-ordered instance stream by tile id -- a single radix pass whose
//       "digit" is the whole tile id (T <= 10240 bins), done here.
// Ties in depth keep ascending Gaussian index through (1)'s stability, exactly like the reference's stable sort, so
// sorted keys / values / ranges are bit-identical.
//
// Kernels (A and B walk the depth-ordered Gaussians warp-cooperatively: one Gaussian at a time per warp, its tiles
// spread over the 32 lanes; all ranking happens in per-warp shared-memory counters -- no global atomics).
// (A variant that flattens the instances of 32 Gaussians over the lanes and ranks with match.any was measured 2.4x
// SLOWER on B200 -- match.any serialises over distinct values -- and was dropped.)
//   A  part_count    per CTA chunk: instances per tile                         -> table[cta][tile]
//   S1 part_colscan  exclusive scan of table over CTAs (per tile), tile totals -> table (in place), tile_total
//   S2 part_tilescan exclusive scan over tiles                                 -> tile_start, ranges, R
//   B  part_scatter  re-walk: rank = tile_start + table[cta][tile] + warp prefix + running count; write point_list
//                    (and the reference-format key) straight to its final slot.
// HBM traffic: ~12 B written per instance (4 B id + 8 B key) + O(CTAs*T) table; nothing is read back per instance.
#include "common.cuh"

namespace fdls {

constexpr int PT_THREADS = 256;
constexpr int PT_WARPS = PT_THREADS / 32;
constexpr int PT_MAX_TILES = 10240;  // 8 warps x 10240 x 2 B + 10240 x 4 B = 200 KB of shared memory

// Work is split by COST (instances + a per-Gaussian overhead), not by Gaussian count: in depth order the near
// (= large on screen) Gaussians come first, so equal Gaussian counts per CTA would leave the first CTAs with most of
// the instances, and equal instance counts would leave the last CTAs with most of the (tiny) Gaussians.  CTA c owns
// the cost span [c*I, (c+1)*I) of the depth-ordered stream (I = ceil(total / ctas), read on the device), warp w an
// eighth of it; a Gaussian whose instances straddle a boundary is shared (its tiles are distinct, so any split
// keeps the per-tile order).
// ctas: see pt_ctas below.
constexpr int PT_SMS = 148;                 // B200
constexpr size_t PT_SMEM_PER_SM = 227 * 1024;
static size_t pt_smem_bytes(int T, bool scatter) {
  const size_t words = (size_t)(T + 1) / 2;
  return (PT_WARPS * words + (scatter ? (size_t)T + 1 : 0)) * sizeof(uint32_t);
}
// Cost units per CTA span: 16 per tile.  Every CTA zeroes, prefixes and reads back 8 warps x T private counters, so
// a span must be worth its counters: at cfg2 (12 M cost units, T = 2040) 740 CTAs of 16 K units spent more
// instructions on their 12 M counters than on the 2.5 M instances; 370 CTAs of 32 K units have the same latency on an
// empty GPU (80 -> 79 us) and give 4280 -> 4950 frames/s with four frames in flight, where the SMs the binning leaves
// alone run another frame's compositor; 40 K units lose again (92 us).  The same ratio is the measured optimum at cfg4
// (T = 5440: 64-96 K units) and cfg5 (T = 8160: 128 K units, 245 -> 279 frames/s against 64 K).  Spans stay below 256 K
// units: CTAs that run together then work on neighbouring stretches of the stream and their 4-byte stores into a
// tile's list share sectors in L2 (round 1: 2.5x faster scatter at R = 185M than one wave of very long spans).
static int pt_ctas(long long capacity, int P, int T) {
  // total cost <= capacity + PT_C * P (PT_C = 32, see below).  A CTA span of cost c holds at most c / 33 Gaussians,
  // which must stay below 65536 (u16 per-tile counters): c <= 2e6.
  const long long cost = capacity + 32LL * P;
  const long long hard = (cost + 1999999) / 2000000;
  long long span = 16LL * T;
  if (span < 16384) span = 16384;
  if (span > 262144) span = 262144;
  long long g = (cost + span - 1) / span;
  // small scenes: at least two CTAs per SM while a span is still worth 4 K units
  const long long floor_g = (cost + 4095) / 4096 < 2 * PT_SMS ? (cost + 4095) / 4096 : 2 * PT_SMS;
  if (g < floor_g) g = floor_g;
  if (g < hard) g = hard;
  if (g < 1) g = 1;
  long long per_sm = (long long)(PT_SMEM_PER_SM / (pt_smem_bytes(T, true) + 1024));
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 8) per_sm = 8;
  const long long wave = per_sm * PT_SMS;  // resident CTAs of the scatter kernel
  if (g > wave) {
    // several waves: a whole number of them (a few CTAs trailing behind a full wave cost a wave's time; cfg4: +3.5 %)
    long long r = (g + wave / 2) / wave * wave;
    if (r < hard) r = (hard + wave - 1) / wave * wave;
    g = r;
  }
  return (int)g;
}

// u16 counters packed two per 32-bit word; returns the previous value of counter t
__device__ __forceinline__ uint32_t cnt_inc(uint32_t* w, uint32_t t) {
  const uint32_t old = atomicAdd(&w[t >> 1], (t & 1u) ? 0x10000u : 1u);
  return (t & 1u) ? (old >> 16) : (old & 0xffffu);
}
__device__ __forceinline__ uint32_t cnt_get(const uint32_t* w, uint32_t t) {
  const uint32_t v = w[t >> 1];
  return (t & 1u) ? (v >> 16) : (v & 0xffffu);
}

// Cost model of the walk: a Gaussian costs its instance count plus PT_C (shuffles, loop set-up and one mostly idle
// inner iteration ~ one full 32-lane iteration).  Spans are equal in COST: W_incl(p) = offsets[p] + PT_C * (p + 1)
// over the V visible Gaussians, which come first in depth order (culled ones carry the key 0xffffffff).
constexpr uint32_t PT_C = 32;
#ifndef PT_GL_
#define PT_GL_ 8
#endif
constexpr int PT_GL = PT_GL_;                 // lanes per Gaussian in the counting walks
constexpr uint32_t PT_SMALL = 8u * PT_GL;    // more tiles than this: the Gaussian takes the whole warp
__device__ __forceinline__ unsigned long long w_incl(const uint32_t* __restrict__ offsets, int p) {
  return (unsigned long long)offsets[p] + (unsigned long long)PT_C * (uint32_t)(p + 1);
}

// smallest position p in [0, V) with W_incl(p) > ws (ws < W_incl(V-1)); 32-ary search, one probe per lane per round
__device__ __forceinline__ int first_pos_above(const uint32_t* __restrict__ offsets, int V, unsigned long long ws,
                                               int lane) {
  int lo = 0, hi = V;
  while (hi - lo > 32) {
    const int step = (hi - lo + 31) / 32;
    const int p = min(lo + (lane + 1) * step, hi) - 1;
    const uint32_t m = __ballot_sync(0xffffffffu, w_incl(offsets, p) > ws);
    const int k = __ffs(m) - 1;  // lane 31 probes hi-1, which is above `ws` by the invariant
    hi = min(lo + (k + 1) * step, hi);
    lo = lo + k * step;
  }
  const uint32_t m = __ballot_sync(0xffffffffu, lo + lane < hi && w_incl(offsets, lo + lane) > ws);
  return lo + __ffs(m) - 1;
}

// Walk the cost span [ws, we) of the depth-ordered stream, one Gaussian at a time per warp, its tiles spread over the
// 32 lanes.  Instance j of the Gaussian at position p sits at cost coordinate excl(p) + PT_C*p + PT_C + j.
// (A variant that flattens the instances of 32 Gaussians over the lanes and ranks with match.any was measured 2.4x
// SLOWER on B200 -- match.any serialises over distinct values -- and was dropped.)
template <bool SCATTER>
__device__ __forceinline__ void walk(unsigned long long ws, unsigned long long we, int V,
                                     const uint32_t* __restrict__ order, const uint32_t* __restrict__ offsets,
                                     const int* __restrict__ radii, const float4* __restrict__ rec, int gx, int gy,
                                     uint32_t* wcnt, const uint32_t* blockbase, uint32_t* __restrict__ point_list,
                                     uint64_t* __restrict__ keys, long long capacity, int lane, int& start) {
  // start: position of the span's first Gaussian; < 0 = not known yet (found by the 32-ary search, four dependent
  // rounds of loads, and handed back so that a second walk of the same span starts at once)
  if (ws >= we) return;
  if (start < 0) start = first_pos_above(offsets, V, ws, lane);
  for (int g0 = start;; g0 += 32) {
    const int pos = g0 + lane;
    uint32_t idx = 0, depth_bits = 0, j0 = 0, j1 = 0, packed = 0, magic = 0;
    bool past = pos >= V;
    if (!past) {
      idx = order[pos];
      const uint32_t incl = offsets[pos];
      const int radius = radii[idx];
      uint32_t count = 0;
      if (radius > 0) {
        const float4 r0 = rec[(size_t)idx * 4];
        if (SCATTER && keys != nullptr) depth_bits = __float_as_uint(rec[(size_t)idx * 4 + 1].z);
        uint2 rmin, rmax;
        get_rect(r0.x, r0.y, radius, gx, gy, rmin, rmax);
        const uint32_t w = rmax.x - rmin.x;
        count = w * (rmax.y - rmin.y);
        packed = rmin.x | (rmin.y << 10) | (w << 20);   // gx, gy < 1024 (partition_supported)
        magic = w > 1 ? 0xffffffffu / w + 1u : 0u;      // floor(i / w) == umulhi(i, magic) while i * w < 2^32
      }
      const unsigned long long w0 = (unsigned long long)(incl - count) + (unsigned long long)PT_C * (uint32_t)pos + PT_C;
      past = w0 - PT_C >= we;
      if (!past) {  // clip the Gaussian's instances [0, count) to this span
        j0 = ws > w0 ? (uint32_t)min((unsigned long long)count, ws - w0) : 0u;
        j1 = we > w0 ? (uint32_t)min((unsigned long long)count, we - w0) : 0u;
      }
    }
    uint32_t todo = __ballot_sync(0xffffffffu, j1 > j0);
    const bool last = __any_sync(0xffffffffu, past);
    if (!SCATTER) {
      // Counting needs no order (the counters are shared-memory atomics): Gaussians with few tiles -- the bulk: ten
      // tiles on average -- go FOUR at a time, eight lanes each, instead of one at a time with two thirds of the warp
      // idle; only the big ones keep the whole warp.
      const uint32_t big = __ballot_sync(0xffffffffu, j1 - j0 > PT_SMALL);
      uint32_t small = todo & ~big;
      todo = big;
      const int grp = lane / PT_GL, sub = lane % PT_GL;
      while (small) {
        int src = -1;
#pragma unroll
        for (int q = 0; q < 32 / PT_GL; ++q) {
          const int b = __ffs(small) - 1;  // -1 once the set is exhausted
          if (q == grp) src = b;
          small &= small - 1;
        }
        const int sl = src < 0 ? lane : src;
        uint32_t b_j0 = __shfl_sync(0xffffffffu, j0, sl);
        uint32_t b_j1 = __shfl_sync(0xffffffffu, j1, sl);
        const uint32_t b_pack = __shfl_sync(0xffffffffu, packed, sl);
        const uint32_t b_magic = __shfl_sync(0xffffffffu, magic, sl);
        if (src < 0) b_j1 = b_j0;
        const uint32_t b_w = b_pack >> 20;
        const uint32_t t0 = ((b_pack >> 10) & 1023u) * gx + (b_pack & 1023u);
        for (uint32_t i = b_j0 + sub; i < b_j1; i += PT_GL) {
          const uint32_t row = b_w > 1 ? __umulhi(i, b_magic) : i;
          cnt_inc(wcnt, t0 + row * gx + (i - row * b_w));
        }
        __syncwarp();
      }
    }
    while (todo) {
      const int src = __ffs(todo) - 1;
      todo &= todo - 1;
      const uint32_t b_j0 = __shfl_sync(0xffffffffu, j0, src);
      const uint32_t b_j1 = __shfl_sync(0xffffffffu, j1, src);
      const uint32_t b_pack = __shfl_sync(0xffffffffu, packed, src);
      const uint32_t b_magic = __shfl_sync(0xffffffffu, magic, src);
      uint32_t b_depth = 0, b_idx = 0;
      if (SCATTER) {
        if (keys != nullptr) b_depth = __shfl_sync(0xffffffffu, depth_bits, src);
        b_idx = __shfl_sync(0xffffffffu, idx, src);
      }
      const uint32_t b_w = b_pack >> 20;
      const uint32_t t0 = ((b_pack >> 10) & 1023u) * gx + (b_pack & 1023u);  // tile of the rect's first cell
      for (uint32_t i = b_j0 + lane; i < b_j1; i += 32) {
        const uint32_t row = b_w > 1 ? __umulhi(i, b_magic) : i;
        const uint32_t t = t0 + row * gx + (i - row * b_w);
        const uint32_t old = cnt_inc(wcnt, t);
        if (SCATTER) {
          const long long slot = (long long)blockbase[t] + old;
          if (slot < capacity) {
            point_list[slot] = b_idx;
            if (keys != nullptr) keys[slot] = ((uint64_t)t << 32) | b_depth;  // reference-format key, on request
          }
        }
      }
      __syncwarp();
    }
    if (last) break;
  }
}

// this warp's cost span; V = visible Gaussians = P - (depth keys whose top byte is 0xff, i.e. the culled ones)
__device__ __forceinline__ void pt_span(const uint32_t* __restrict__ num_rendered, const uint32_t* __restrict__ depth_hist,
                                        int P, int warp, unsigned long long& ws, unsigned long long& we, int& V) {
  const uint32_t R = num_rendered[0];
  V = P - (int)depth_hist[3 * 256 + 255];
  const unsigned long long total = R ? (unsigned long long)R + (unsigned long long)PT_C * (uint32_t)V : 0ull;
  const unsigned long long I = (total + gridDim.x - 1) / gridDim.x;
  const unsigned long long Iw = (I + PT_WARPS - 1) / PT_WARPS;
  const unsigned long long c0 = (unsigned long long)blockIdx.x * I;
  ws = c0 + (unsigned long long)warp * Iw;
  we = ws + Iw;
  if (we > c0 + I) we = c0 + I;
  if (we > total) we = total;
  if (ws > total) ws = total;
}

__global__ void __launch_bounds__(PT_THREADS)
part_count_kernel(int P, int T, const uint32_t* __restrict__ num_rendered,
                  const uint32_t* __restrict__ depth_hist, const uint32_t* __restrict__ order,
                  const uint32_t* __restrict__ offsets, const int* __restrict__ radii, const float4* __restrict__ rec,
                  int gx, int gy, uint32_t* __restrict__ table) {
  pdl_grid_sync();
  // ONE counter set for the whole CTA: only the CTA's totals are wanted here and counting needs no order, so the
  // warps' atomics may interleave (the scatter kernel needs a set per warp).  T / 2 words instead of 4 T: at cfg5
  // (T = 8160) 16 KB per CTA instead of 131 KB, i.e. the SM holds eight CTAs instead of one.  A CTA span holds fewer
  // than 65536 Gaussians (pt_ctas), so the u16 counters cannot overflow.
  extern __shared__ uint32_t pt_smem[];
  const int words = (T + 1) / 2;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < words; i += PT_THREADS) pt_smem[i] = 0;
  __syncthreads();
  unsigned long long ws, we;
  int V;
  pt_span(num_rendered, depth_hist, P, warp, ws, we, V);
  int start = -1;
  walk<false>(ws, we, V, order, offsets, radii, rec, gx, gy, pt_smem, nullptr, nullptr, nullptr, 0, lane, start);
  __syncthreads();
  uint32_t* row = table + (size_t)blockIdx.x * T;
  for (int t = tid; t < T; t += PT_THREADS) row[t] = cnt_get(pt_smem, t);
}

// per tile: exclusive scan over the CTAs (table is [ctas][T], coalesced along t); tile_total[t] = column sum.
// Each thread sums its slice of rows, the slice sums of a tile are scanned through shared memory, then the slice is
// rewritten as running prefixes (second read served by L1/L2).
__device__ __forceinline__ void tilescan_body(int T, const uint32_t* tile_total, uint32_t* __restrict__ tile_start,
                                              uint2* __restrict__ ranges, uint32_t* __restrict__ num_rendered,
                                              long long capacity);

// The LAST block to finish (ticket in num_rendered[3], zeroed by the preprocess kernel and reset here) goes on to scan
// the tile totals (tilescan_body), which saves the one-block launch that used to follow.
__global__ void __launch_bounds__(1024)
part_colscan_kernel(int T, int ctas, uint32_t* __restrict__ table, uint32_t* __restrict__ tile_total,
                    uint32_t* __restrict__ tile_start, uint2* __restrict__ ranges,
                    uint32_t* __restrict__ num_rendered, long long capacity) {
  pdl_grid_sync();
  // CTA = 8 tiles x 128 row groups (a warp reads 4 row groups x 32 bytes: whole sectors); T/8 CTAs fill the GPU
  __shared__ uint32_t s_sum[128][9];
  __shared__ uint32_t s_last;
  const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
  const int t = blockIdx.x * 8 + tx;
  const int per = (ctas + 127) / 128;
  const int r0 = min(ty * per, ctas), r1 = min(r0 + per, ctas);
  uint32_t sum = 0;
  if (t < T)
    for (int r = r0; r < r1; ++r) sum += table[(size_t)r * T + t];
  s_sum[ty][tx] = sum;
  __syncthreads();
  uint32_t run = 0;
  for (int k = 0; k < ty; ++k) run += s_sum[k][tx];
  if (t < T) {
    for (int r = r0; r < r1; ++r) {
      const uint32_t v = table[(size_t)r * T + t];
      table[(size_t)r * T + t] = run;
      run += v;
    }
    if (ty == 127) tile_total[t] = run;  // the last slice ends at the column total (empty slices pass `run` through)
  }
  __threadfence();  // every writer publishes its tile totals before the block takes its ticket
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(&num_rendered[3], 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  tilescan_body(T, tile_total, tile_start, ranges, num_rendered, capacity);
  if (threadIdx.x == 0) num_rendered[3] = 0;
}

__global__ void __launch_bounds__(PT_THREADS)
part_scatter_kernel(int P, int T, const uint32_t* __restrict__ num_rendered,
                    const uint32_t* __restrict__ depth_hist, const uint32_t* __restrict__ order,
                    const uint32_t* __restrict__ offsets, const int* __restrict__ radii,
                    const float4* __restrict__ rec, int gx, int gy, const uint32_t* __restrict__ table,
                    const uint32_t* __restrict__ tile_start, uint32_t* __restrict__ point_list,
                    uint64_t* __restrict__ keys, long long capacity) {
  pdl_grid_sync();
  extern __shared__ uint32_t pt_smem[];
  const int words = (T + 1) / 2;
  uint32_t* blockbase = pt_smem + PT_WARPS * words;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < PT_WARPS * words; i += PT_THREADS) pt_smem[i] = 0;
  __syncthreads();
  unsigned long long ws, we;
  int V;
  pt_span(num_rendered, depth_hist, P, warp, ws, we, V);
  // walk 1: per-warp tile counts (same as kernel A)
  int start = -1;
  walk<false>(ws, we, V, order, offsets, radii, rec, gx, gy, pt_smem + warp * words, nullptr, nullptr, nullptr, 0, lane,
              start);
  __syncthreads();
  // counts -> exclusive prefix across the warps of this CTA (in place); global base of the CTA per tile
  const uint32_t* row = table + (size_t)blockIdx.x * T;
  for (int w2 = tid; w2 < words; w2 += PT_THREADS) {  // two tiles per word; the two u16 lanes scan independently
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < PT_WARPS; ++w) {
      const uint32_t c = pt_smem[w * words + w2];
      pt_smem[w * words + w2] = run;
      run += c;  // no carry between halves: a CTA span holds < 65536 Gaussians (pt_ctas), hence < 65536 per tile
    }
    const int t0 = 2 * w2;
    blockbase[t0] = tile_start[t0] + row[t0];
    if (t0 + 1 < T) blockbase[t0 + 1] = tile_start[t0 + 1] + row[t0 + 1];
  }
  __syncthreads();
  // walk 2: the running counters now start at each warp's prefix -> final slots
  walk<true>(ws, we, V, order, offsets, radii, rec, gx, gy, pt_smem + warp * words, blockbase, point_list, keys,
             capacity, lane, start);
}

// tile_scan_kernel lives in binning_scan below: exclusive scan over tile totals -> tile_start, ranges, R
__device__ __forceinline__ void tilescan_body(int T, const uint32_t* tile_total, uint32_t* __restrict__ tile_start,
                                              uint2* __restrict__ ranges, uint32_t* __restrict__ num_rendered,
                                              long long capacity) {
  __shared__ uint32_t s_warp[33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t carry = 0;
  for (int base = 0; base < T; base += 1024) {
    const int t = base + threadIdx.x;
    const uint32_t c = t < T ? __ldcg(&tile_total[t]) : 0;
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
      // the reference leaves empty tiles at (0,0) (cudaMemset, rasterizer_impl.cu:314); clip to the workspace so that
      // the compositor never walks past it after a capacity overflow (the caller then re-runs with a larger one)
      const long long lo = min((long long)start, capacity), hi = min((long long)start + c, capacity);
      ranges[t] = hi > lo ? make_uint2((uint32_t)lo, (uint32_t)hi) : make_uint2(0u, 0u);
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

__global__ void __launch_bounds__(1024)
part_tilescan_kernel(int T, const uint32_t* __restrict__ tile_total, uint32_t* __restrict__ tile_start,
                     uint2* __restrict__ ranges, uint32_t* __restrict__ num_rendered, long long capacity) {
  tilescan_body(T, tile_total, tile_start, ranges, num_rendered, capacity);
}


bool partition_supported(int gx, int gy) { return gx < 1024 && gy < 1024 && gx * gy <= PT_MAX_TILES; }
int partition_ctas(long long capacity, int P, int T) { return pt_ctas(capacity < 1 ? 1 : capacity, P < 1 ? 1 : P, T); }
size_t partition_table_bytes(long long capacity, int P, int T) {
  return align_up((size_t)partition_ctas(capacity, P, T) * T * sizeof(uint32_t));
}

// count -> column scan -> tile scan -> scatter, all on `st`; R is read from num_rendered[0] on the device.
cudaError_t launch_partition(int P, int gx, int gy, const uint32_t* order, const uint32_t* offsets, const int* radii,
                             const float4* rec, uint32_t* table, uint32_t* tile_total, uint32_t* tile_start,
                             uint2* ranges, uint32_t* num_rendered, uint32_t* point_list, uint64_t* keys,
                             const uint32_t* depth_hist, long long capacity, cudaStream_t st) {
  const int T = gx * gy, ctas = partition_ctas(capacity, P, T);
  static SmemOptIn optin_a, optin_b;  // per device (common.cuh)
  const size_t sa = (size_t)((T + 1) / 2) * sizeof(uint32_t), sb = pt_smem_bytes(T, true);
  if (cudaError_t e = optin_a.ensure(part_count_kernel, sa)) return e;
  if (cudaError_t e = optin_b.ensure(part_scatter_kernel, sb)) return e;
  launch_pdl(part_count_kernel, dim3(ctas), dim3(PT_THREADS), sa, st, P, T, num_rendered, depth_hist, order, offsets, radii, rec, gx, gy, table);
  count_launch();
  launch_pdl(part_colscan_kernel, dim3(cdiv(T, 8)), dim3(1024), 0, st, T, ctas, table, tile_total, tile_start, ranges, num_rendered, capacity);
  count_launch();
  launch_pdl(part_scatter_kernel, dim3(ctas), dim3(PT_THREADS), sb, st, P, T, num_rendered, depth_hist, order, offsets, radii, rec, gx, gy, table, tile_start, point_list, keys, capacity);
  count_launch();
  return cudaGetLastError();
}

}  // namespace fdls
