// radix_sort.cu -- stable LSD "onesweep" radix sort of (u64 key, u32 value) pairs on key bits [begin_bit, end_bit).
//
// Replaces cub::DeviceRadixSort::SortPairs as called at rasterizer_impl.cu:307-312 on the (tile << 32 | depth bits)
// keys.  Written from the published onesweep scheme (one upfront multi-pass histogram, then ONE kernel per 8-bit
// digit that ranks a tile of keys in shared memory and resolves its global offsets by decoupled look-back over the
// per-tile digit counts of the preceding tiles), sized for B200:
//   * 256 threads x IPT keys per CTA; IPT is chosen from the problem size so that even the 300k-element depth sort
//     fills the 148 SMs (IPT=4 -> 1024-key tiles) while the multi-million-instance sort amortises its per-CTA scan
//     and look-back (IPT=8);
//   * one shared exchange buffer is used twice (keys, then values) and values are loaded only after ranking, which
//     keeps the CTA at <= 64 registers/thread and ~25 KB of shared memory -> 4-5 CTAs per SM;
//   * the element count is read from DEVICE memory (no host sync on num_rendered); CTAs past the end exit;
//   * tiles are handed out by an atomic ticket so a CTA only ever waits on CTAs that already started;
//   * a caller that generates the keys itself (duplicate_keys_kernel) can accumulate the digit histograms on the fly
//     and skip the histogram pass (sort_prepare / sort_hist / sort_run).
// Stability: items are ranked in (warp, item, lane) order == global index order; equal digits keep that order.
#include "common.cuh"

namespace fdls {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_BINS = 256;
constexpr int RS_MAX_PASSES = 8;
constexpr uint32_t FLAG_AGG = 1u << 30, FLAG_PREFIX = 2u << 30, FLAG_MASK = 3u << 30, VAL_MASK = ~FLAG_MASK;
constexpr unsigned FULL = 0xffffffffu;

template <int IPT>
struct RsSmem {
  union {
    uint64_t keys[RS_THREADS * IPT];
    uint32_t vals[RS_THREADS * IPT];
  } x;
  uint32_t warp_hist[RS_WARPS][RS_BINS + 8];
  uint32_t digit_start[RS_BINS];
  uint32_t global_base[RS_BINS];
  uint32_t scan_warp[40];
  uint32_t tile;
};

__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}
__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Lanes holding the same 9-bit digit (bin 256 = padding).  Nine ballots: a fixed ~30-instruction cost, where the
// hardware match.any serialises over the distinct values in the warp (measured slower on B200 for 8-bit digits).
__device__ __forceinline__ uint32_t rs_match9(uint32_t d) {
  uint32_t peers = FULL;
#pragma unroll
  for (int b = 0; b < 9; ++b) {
    const bool bit = (d >> b) & 1u;
    const uint32_t bal = __ballot_sync(FULL, bit);
    peers &= bit ? bal : ~bal;
  }
  return peers;
}

__device__ __forceinline__ uint32_t rs_warp_incl(uint32_t v, int lane) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t t = __shfl_up_sync(FULL, v, d);
    if (lane >= d) v += t;
  }
  return v;
}
// exclusive scan of one value per thread over the 256-thread CTA
__device__ __forceinline__ uint32_t rs_block_excl(uint32_t v, uint32_t* s_warp) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = rs_warp_incl(v, lane);
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < RS_WARPS ? s_warp[lane] : 0;
    uint32_t wi = rs_warp_incl(w, lane);
    s_warp[lane] = wi - w;
  }
  __syncthreads();
  uint32_t r = incl - v + s_warp[warp];
  __syncthreads();
  return r;
}

// Upfront histogram of every pass's digit (reads the keys once).
__global__ void __launch_bounds__(RS_THREADS)
rs_histogram_kernel(const uint64_t* __restrict__ keys, long long n_max, const uint32_t* __restrict__ n_dev, int passes,
                    int begin_bit, int end_bit, uint32_t* __restrict__ global_hist) {
  pdl_grid_sync();
  constexpr int HT = RS_THREADS * 8;
  __shared__ uint32_t s_hist[RS_MAX_PASSES][RS_BINS];
  const long long n = n_dev ? min((long long)n_dev[0], n_max) : n_max;
  const long long base = (long long)blockIdx.x * HT;
  if (base >= n) return;
  for (int i = threadIdx.x; i < passes * RS_BINS; i += RS_THREADS) (&s_hist[0][0])[i] = 0;
  __syncthreads();
  const uint32_t lt = lanemask_lt();
#pragma unroll 4
  for (int i = 0; i < 8; ++i) {
    const long long k = base + i * RS_THREADS + threadIdx.x;
    const bool valid = k < n;
    const uint64_t key = valid ? keys[k] : 0;
    const unsigned act = __ballot_sync(FULL, valid);
    if (valid) {
      for (int p = 0; p < passes; ++p) {
        const int shift = begin_bit + p * 8;
        const uint32_t mask = (1u << min(8, end_bit - shift)) - 1;
        const uint32_t d = (uint32_t)(key >> shift) & mask;
        atomicAdd(&s_hist[p][d], 1u);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * RS_BINS; i += RS_THREADS) {
    uint32_t c = (&s_hist[0][0])[i];
    if (c) atomicAdd(&global_hist[i], c);
  }
}

template <int IPT>
__global__ void __launch_bounds__(RS_THREADS)
rs_onesweep_kernel(const uint64_t* __restrict__ keys_in, uint64_t* __restrict__ keys_out,
                   const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out, long long n_max,
                   const uint32_t* __restrict__ n_dev, int shift, int bits, const uint32_t* __restrict__ global_hist,
                   uint32_t* __restrict__ tile_state, uint32_t* __restrict__ ticket, uint32_t* __restrict__ error_flag) {
  pdl_grid_sync();
  constexpr int TILE_KEYS = RS_THREADS * IPT;
  extern __shared__ __align__(16) unsigned char rs_smem_raw[];
  RsSmem<IPT>& s = *reinterpret_cast<RsSmem<IPT>*>(rs_smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s.tile = atomicAdd(ticket, 1u);
  for (int i = tid; i < RS_WARPS * (RS_BINS + 8); i += RS_THREADS) (&s.warp_hist[0][0])[i] = 0;
  __syncthreads();
  const uint32_t tile = s.tile;
  const long long n = n_dev ? min((long long)n_dev[0], n_max) : n_max;
  const long long tile_base = (long long)tile * TILE_KEYS;
  if (tile_base >= n) return;
  const int valid_count = (int)min((long long)TILE_KEYS, n - tile_base);
  const uint32_t mask = (1u << bits) - 1;

  // ---- load keys (warp-striped) and compute every item's peer set up front (independent match.any's)
  uint64_t key[IPT];
  unsigned peers[IPT];
  uint16_t rank[IPT];
  const int warp_base = warp * 32 * IPT;
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const int li = warp_base + i * 32 + lane;
    key[i] = (li < valid_count) ? keys_in[tile_base + li] : ~0ull;
  }
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const int li = warp_base + i * 32 + lane;
    const uint32_t d = (li < valid_count) ? ((uint32_t)(key[i] >> shift) & mask) : RS_BINS;  // bin 256 = padding
    peers[i] = rs_match9(d);
  }
  // ---- rank inside the warp: running per-digit counters live in this warp's private shared histogram
  const uint32_t lt = lanemask_lt();
  uint32_t* wh = s.warp_hist[warp];
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const int li = warp_base + i * 32 + lane;
    const uint32_t d = (li < valid_count) ? ((uint32_t)(key[i] >> shift) & mask) : RS_BINS;
    const uint32_t old = wh[d];  // every lane of a peer group reads the same counter (broadcast) ...
    __syncwarp();
    if ((peers[i] & lt) == 0) wh[d] = old + __popc(peers[i]);  // ... then its leader bumps it
    __syncwarp();
    rank[i] = (uint16_t)(old + __popc(peers[i] & lt));
  }
  __syncthreads();

  // ---- per-digit: exclusive scan across warps, CTA count
  const int d = tid;  // one digit per thread (RS_THREADS == RS_BINS)
  uint32_t count = 0;
#pragma unroll
  for (int w = 0; w < RS_WARPS; ++w) {
    uint32_t c = s.warp_hist[w][d];
    s.warp_hist[w][d] = count;
    count += c;
  }
  // ---- publish the CTA's digit count, then look back for the exclusive prefix over preceding tiles
  uint32_t* my_state = tile_state + (size_t)tile * RS_BINS + d;
  st_relaxed(my_state, (tile == 0 ? FLAG_PREFIX : FLAG_AGG) | count);
  uint32_t excl = 0;
  if (tile > 0) {
    // Look-back, 8 predecessors per step: the 8 loads are in flight together (one L2 round trip instead of eight),
    // then consumed nearest-first; a not-yet-published entry ends the step and is polled again.
    constexpr int LB = 8;
    long long t = (long long)tile - 1;
    uint32_t spins = 0;
    bool done = false;
    while (!done) {
      uint32_t v[LB];
#pragma unroll
      for (int i = 0; i < LB; ++i) v[i] = (t - i >= 0) ? ld_relaxed(tile_state + (size_t)(t - i) * RS_BINS + d) : FLAG_PREFIX;
#pragma unroll
      for (int i = 0; i < LB; ++i) {
        if (done) break;
        const uint32_t f = v[i] & FLAG_MASK;
        if (f == 0) {  // predecessor t-i has not published yet: resume from it
          t -= i;
          // Forward progress: tiles are handed out by ticket, so every predecessor belongs to a CTA that is already
          // running and never waits on a later tile -- it will publish.  The poll limit (>= 2^26 x 100 ns = 6.7 s of
          // waiting on ONE entry) only exists so that a corrupted scratch cannot hang the GPU; reaching it is never
          // silent: the flag is set and the kernel traps, so the stream's next CUDA call reports a launch failure
          // instead of handing out a mis-sorted frame.
          if (++spins > (1u << 26)) {
            atomicExch(error_flag, 1u);
            __trap();
          } else if (i == 0) {
            __nanosleep(spins < 64 ? 20 : 100);
          }
          break;
        }
        excl += v[i] & VAL_MASK;
        if (f == FLAG_PREFIX) done = true;
        if (i == LB - 1 && !done) t -= LB;
      }
    }
    st_relaxed(my_state, FLAG_PREFIX | ((excl + count) & VAL_MASK));
  }
  // ---- local digit starts and global bases
  const uint32_t local_start = rs_block_excl(count, s.scan_warp);
  const uint32_t ghist_excl = rs_block_excl(global_hist[d], s.scan_warp);
  s.digit_start[d] = local_start;
  s.global_base[d] = ghist_excl + excl - local_start;
  __syncthreads();

  // ---- keys: exchange through shared memory so equal digits become contiguous, then scatter
  uint16_t pos[IPT];
  uint32_t val[IPT];
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const int li = warp_base + i * 32 + lane;
    pos[i] = 0;
    val[i] = 0;
    if (li < valid_count) {
      const uint32_t dg = (uint32_t)(key[i] >> shift) & mask;
      pos[i] = (uint16_t)(s.digit_start[dg] + wh[dg] + rank[i]);
      s.x.keys[pos[i]] = key[i];
      val[i] = vals_in[tile_base + li];  // in flight while the keys are scattered
    }
  }
  __syncthreads();
  uint32_t dst[IPT];
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const int p = i * RS_THREADS + tid;
    dst[i] = 0;
    if (p < valid_count) {
      const uint64_t k = s.x.keys[p];
      const uint32_t dg = (uint32_t)(k >> shift) & mask;
      dst[i] = s.global_base[dg] + p;
      keys_out[dst[i]] = k;
    }
  }
  __syncthreads();
  // ---- values: same exchange buffer, same positions
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const int li = warp_base + i * 32 + lane;
    if (li < valid_count) s.x.vals[pos[i]] = val[i];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const int p = i * RS_THREADS + tid;
    if (p < valid_count) vals_out[dst[i]] = s.x.vals[p];
  }
}

// ---------------------------------------------------------------------------------------------------------------
struct RsScratch {
  uint32_t* global_hist;  // [RS_MAX_PASSES][256]
  uint32_t* tickets;      // [RS_MAX_PASSES]
  uint32_t* error_flag;   // [1]
  uint32_t* tile_state;   // [passes][tiles][256]
  size_t total_bytes;
};

static int rs_ipt(long long n_max) { return n_max <= (1 << 20) ? 4 : 8; }

static RsScratch rs_carve(void* scratch, long long n_max, int passes) {
  RsScratch r;
  char* p = (char*)scratch;
  r.global_hist = (uint32_t*)p;
  p += RS_MAX_PASSES * RS_BINS * sizeof(uint32_t);
  r.tickets = (uint32_t*)p;
  p += 16 * sizeof(uint32_t);
  r.error_flag = (uint32_t*)p;
  p += 16 * sizeof(uint32_t);
  r.tile_state = (uint32_t*)p;
  const size_t tiles = (size_t)cdiv(n_max, RS_THREADS * rs_ipt(n_max));
  p += (size_t)passes * tiles * RS_BINS * sizeof(uint32_t);
  r.total_bytes = align_up((size_t)(p - (char*)scratch));
  return r;
}

size_t sort_scratch_bytes(long long n_max) { return rs_carve(nullptr, n_max < 1 ? 1 : n_max, RS_MAX_PASSES).total_bytes; }

// Zero the scratch (histograms, tickets, tile states).  Must precede sort_run and any external histogram accumulation.
// zero_before: bytes directly in front of the scratch that the same memset clears (the scan's state words, api.cu)
int sort_prepare(long long n_max, int begin_bit, int end_bit, void* scratch, size_t scratch_bytes, cudaStream_t st,
                 size_t zero_before) {
  if (begin_bit < 0 || end_bit <= begin_bit || end_bit > 64) return FDLS_EINVAL;
  if (n_max <= 0) return FDLS_OK;
  const int passes = (end_bit - begin_bit + 7) / 8;
  RsScratch sc = rs_carve(scratch, n_max, passes);
  if (scratch_bytes < sc.total_bytes) return FDLS_EWORKSPACE;
  return cudaMemsetAsync((char*)scratch - zero_before, 0, sc.total_bytes + zero_before, st) == cudaSuccess ? FDLS_OK
                                                                                                             : FDLS_ECUDA;
}

// [pass][256] digit counters inside the scratch (for callers that build the histogram while generating the keys)
uint32_t* sort_hist(void* scratch) { return (uint32_t*)scratch; }

template <int IPT>
static int rs_launch_passes(uint64_t* kin, uint64_t* kout, uint32_t* vin, uint32_t* vout, long long n_max,
                            const uint32_t* n_dev, int begin_bit, int end_bit, int passes, RsScratch& sc,
                            cudaStream_t st) {
  static SmemOptIn optin;  // per device (common.cuh); one per IPT instantiation
  if (optin.ensure(rs_onesweep_kernel<IPT>, sizeof(RsSmem<IPT>)) != cudaSuccess) return FDLS_ECUDA;
  const int tiles = cdiv(n_max, RS_THREADS * IPT);
  for (int p = 0; p < passes; ++p) {
    const int shift = begin_bit + p * 8;
    const int bits = end_bit - shift < 8 ? end_bit - shift : 8;
    launch_pdl(rs_onesweep_kernel<IPT>, dim3(tiles), dim3(RS_THREADS), sizeof(RsSmem<IPT>), st, kin, kout, vin, vout, n_max, n_dev, shift, bits, sc.global_hist + p * RS_BINS, sc.tile_state + (size_t)p * tiles * RS_BINS, sc.tickets + p, sc.error_flag);
    count_launch();
    uint64_t* tk = kin; kin = kout; kout = tk;
    uint32_t* tv = vin; vin = vout; vout = tv;
  }
  return FDLS_OK;
}

// returns selector (0: result in a, 1: result in b) or a negative FDLS error.  have_hist: the caller already
// accumulated the per-pass digit histograms into sort_hist(scratch) after sort_prepare().
int sort_run(uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, long long n_max,
             const uint32_t* n_dev, int begin_bit, int end_bit, void* scratch, size_t scratch_bytes, bool have_hist,
             cudaStream_t st) {
  if (begin_bit < 0 || end_bit <= begin_bit || end_bit > 64) return FDLS_EINVAL;
  const int passes = (end_bit - begin_bit + 7) / 8;
  if (n_max <= 0) return passes & 1;
  RsScratch sc = rs_carve(scratch, n_max, passes);
  if (scratch_bytes < sc.total_bytes) return FDLS_EWORKSPACE;
  if (!have_hist) {
    launch_pdl(rs_histogram_kernel, dim3(cdiv(n_max, RS_THREADS * 8)), dim3(RS_THREADS), 0, st, keys_a, n_max, n_dev, passes, begin_bit, end_bit, sc.global_hist);
    count_launch();
  }
  int rc = rs_ipt(n_max) == 4
               ? rs_launch_passes<4>(keys_a, keys_b, vals_a, vals_b, n_max, n_dev, begin_bit, end_bit, passes, sc, st)
               : rs_launch_passes<8>(keys_a, keys_b, vals_a, vals_b, n_max, n_dev, begin_bit, end_bit, passes, sc, st);
  if (rc < 0) return rc;
  if (cudaGetLastError() != cudaSuccess) return FDLS_ECUDA;
  return passes & 1;
}

int sort_pairs(uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, long long n_max,
               const uint32_t* n_dev, int begin_bit, int end_bit, void* scratch, size_t scratch_bytes,
               cudaStream_t st) {
  int rc = sort_prepare(n_max, begin_bit, end_bit, scratch, scratch_bytes, st, 0);
  if (rc < 0) return rc;
  return sort_run(keys_a, keys_b, vals_a, vals_b, n_max, n_dev, begin_bit, end_bit, scratch, scratch_bytes, false, st);
}

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free variant (fdls_sort_pairs_static): the scratch of sort_pairs lives in a file-scope device array instead
// of a caller-provided workspace, for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that
// forbid cudaMalloc).  Every GPU that loads the library holds the array (8.0 MiB) until exit.  All calls on a device
// share it: calls that overlap on the GPU (different streams) race.  sort_prepare zeroes the part a call uses, so the
// previous call's contents never leak into the next.
constexpr size_t RS_STATIC_BYTES =
    ((size_t)RS_MAX_PASSES * RS_BINS * sizeof(uint32_t) + 32 * sizeof(uint32_t) +
     (size_t)RS_MAX_PASSES * (FDLS_SORT_STATIC_MAX_PAIRS / (RS_THREADS * 8)) * RS_BINS * sizeof(uint32_t) + 255) / 256 * 256;
static __device__ __align__(256) unsigned char g_rs_static_scratch[RS_STATIC_BYTES];

// device address of the array on the CURRENT device (the symbol's name is a host-side stand-in, not a device pointer)
static int rs_static_scratch(void** out) {
  constexpr int MAX_DEVICES = 64;
  static std::mutex mu;
  static void* ptrs[MAX_DEVICES] = {};
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) return FDLS_ECUDA;
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    if (cudaGetSymbolAddress(&p, g_rs_static_scratch) != cudaSuccess || !p) return FDLS_ECUDA;
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return FDLS_OK;
}

int sort_pairs_static(uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, long long n_max,
                      const uint32_t* n_dev, int begin_bit, int end_bit, cudaStream_t st) {
  if (n_max > FDLS_SORT_STATIC_MAX_PAIRS) return FDLS_ELIMIT;
  if (sort_scratch_bytes(n_max) > RS_STATIC_BYTES) return FDLS_ELIMIT;  // cannot happen below the limit
  void* scratch = nullptr;
  int rc = rs_static_scratch(&scratch);
  if (rc < 0) return rc;
  return sort_pairs(keys_a, keys_b, vals_a, vals_b, n_max, n_dev, begin_bit, end_bit, scratch, RS_STATIC_BYTES, st);
}

}  // namespace fdls
