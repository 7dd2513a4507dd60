// deform_mlp.cu -- the MLP half of the deformation network (scene/deformation.py:45-182) on the 5th-generation tensor
// cores: tcgen05.mma (kind::f16, fp16 x 2 operand split => fp32-grade products, see umma.cuh), accumulators in TMEM,
// activations never leave the SM between layers.
//
// Part 1 (this file, first): ONE linear layer Y = act(X W^T + b) as a stand-alone kernel -- the probe that pins the
// descriptor conventions against torch on the GPU box (tests/test_gpu_deform.py) and the building block of the fused
// network below.
#include "common.cuh"
#include "umma.cuh"

namespace fdls {

using namespace umma;

constexpr int DM_ROWS = 128;  // points per CTA tile = TMEM lanes = UMMA M

// the three fp16 x fp16 terms of an fp32-grade product, smallest first: (part of A, part of B)
constexpr int DM_NTERMS = 3;
__device__ constexpr int DM_TERMS[DM_NTERMS][2] = {{1, 0}, {0, 1}, {0, 0}};

// Issue D[128 x N] (+)= A[128 x K] * B[N x K]^T over `ksteps` K=16 steps.  A, B: three parts each, `a_part` / `b_part`
// bytes apart, canonical slabs (umma.cuh).  One thread.
__device__ __forceinline__ void dm_issue_gemm(uint32_t tmem_d, uint32_t a_base, uint32_t a_part, uint32_t b_base,
                                              uint32_t b_part, int N, int ksteps, bool accumulate_first) {
  const uint32_t idesc = instr_desc_f16(DM_ROWS, N);
  const uint32_t a_slab = DM_ROWS * 16, b_slab = (uint32_t)N * 16;
  bool acc = accumulate_first;
  for (int ks = 0; ks < ksteps; ++ks) {
#pragma unroll
    for (int t = 0; t < DM_NTERMS; ++t) {
      const uint64_t da = smem_desc(a_base + DM_TERMS[t][0] * a_part + ks * 2 * a_slab, a_slab, 128);
      const uint64_t db = smem_desc(b_base + DM_TERMS[t][1] * b_part + ks * 2 * b_slab, b_slab, 128);
      mma_f16(tmem_d, da, db, idesc, acc);
      acc = true;
    }
  }
}

// ---- probe / building block: Y[rows x N] = act(X[rows x K] W[N x K]^T + b), K % 16 == 0, N % 8 == 0, N <= 128, K <= 128
__global__ void __launch_bounds__(DM_ROWS)
umma_linear_kernel(int rows, int N, int K, const float* __restrict__ X, const float* __restrict__ W,
                   const float* __restrict__ bias, float* __restrict__ Y, int relu) {
  extern __shared__ __align__(128) unsigned char dm_smem[];
  __shared__ uint64_t s_bar;
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t a_part = (uint32_t)(K / 8) * DM_ROWS * 16;  // bytes of one fp16 part of the activation tile
  const uint32_t b_part = (uint32_t)(K / 8) * N * 16;
  char* sA = reinterpret_cast<char*>(dm_smem);
  char* sB = sA + PARTS * a_part;

  if (warp == 0) tmem_alloc<128>(&s_tmem);
  if (tid == 0) {
    mbar_init(&s_bar, 1);
    fence_mbar_init();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = s_tmem;

  // weights: row n, 8 k at a time
  for (int i = tid; i < N * (K / 8); i += DM_ROWS) {
    const int n = i % N, slab = i / N;
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = W[(size_t)n * K + slab * 8 + e];
    store_row8_split(sB + (size_t)slab * N * 16 + n * 16, b_part, v);
  }
  // activations: thread = row
  const int row = blockIdx.x * DM_ROWS + tid;
  for (int slab = 0; slab < K / 8; ++slab) {
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = row < rows ? X[(size_t)row * K + slab * 8 + e] : 0.f;
    store_row8_split(sA + (size_t)slab * DM_ROWS * 16 + tid * 16, a_part, v);
  }
  fence_proxy_async();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after_sync();
    dm_issue_gemm(tmem, smem_u32(sA), a_part, smem_u32(sB), b_part, N, K / 16, false);
    mma_commit(&s_bar);
  }
  mbar_wait(&s_bar, 0);
  tc_fence_after_sync();
  for (int c = 0; c < N; c += 8) {
    float v[8];
    tmem_ld8(tmem + ((uint32_t)(warp * 32) << 16) + c, v);
    if (row < rows) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        float y = v[e] + (bias ? bias[c + e] : 0.f);
        if (relu) y = fmaxf(y, 0.f);
        Y[(size_t)row * N + c + e] = y;
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<128>(tmem);
}

int umma_linear(int rows, int N, int K, const float* X, const float* W, const float* bias, float* Y, int relu,
                cudaStream_t st) {
  if (rows < 0 || N < 8 || N > 128 || (N % 8) || K < 16 || K > 128 || (K % 16)) return FDLS_EINVAL;
  if (rows == 0) return FDLS_OK;
  const size_t smem = (size_t)PARTS * (K / 8) * 16 * (DM_ROWS + N);
  static SmemOptIn optin;
  if (optin.ensure(umma_linear_kernel, smem) != cudaSuccess) return FDLS_ECUDA;
  count_launch();
  umma_linear_kernel<<<cdiv(rows, DM_ROWS), DM_ROWS, smem, st>>>(rows, N, K, X, W, bias, Y, relu);
  return cudaGetLastError() == cudaSuccess ? FDLS_OK : FDLS_ECUDA;
}


// ======================================================================================================================
// Part 2: the fused network.
//
// Per 128-point tile (thread = point = TMEM lane):
//   X tile -> fp16x2 slabs (sX)                                   trunk GEMM  X W0^T          -> TMEM cols [0, W)
//   for every head:  ReLU(trunk + b0) -> sA                       layer-1 GEMM sA W1^T        -> TMEM cols [W, 2W)
//                    ReLU(hidden + b1) -> sA                      layer-2 GEMM sA W2^T        -> TMEM cols [W, W + Npad)
//                    act(base + out + b2) -> global
// The trunk accumulator stays in TMEM for the whole tile and is re-read per head (shared memory holds ONE activation
// buffer: 3 parts x 32 KB at W = 128).  Weights arrive as a fixed cyclic sequence of <= 24 KB operand chunks through a
// five-slot TMA ring (cp.async.bulk): while the tensor core works on chunk g the copies of chunks g+1 .. g+3 are in flight (an L2
// round trip is longer than a chunk's MMAs), and slot (g+3) % 5 is free because chunk g-2's MMAs have completed
// (mbarrier of its tcgen05.commit).  The input tile aliases the activation buffer.
// ======================================================================================================================
constexpr int DM_CHUNK_BYTES = 16384;  // 4 slabs x 128 rows x 16 B x PARTS
constexpr int DM_MAX_CHUNKS = 2 + FDLS_DEFORM_MAX_HEADS * (4 + 3);
constexpr int DM_KIND_TRUNK = 0, DM_KIND_HIDDEN = 1, DM_KIND_OUT = 2;

struct DmChunk {
  uint32_t off;     // byte offset of the chunk in the prepared buffer
  uint32_t bytes;   // PARTS x slabs x N x 16
  uint16_t N;       // rows of the B operand (= output columns, padded)
  uint16_t slab0;   // first A slab (8 k each) this chunk multiplies
  uint16_t slabs;   // slabs in this chunk (even)
  uint8_t kind, head;
  uint8_t first, last;  // first / last chunk of its GEMM
};
struct DmPlan {
  int n_chunks;
  int npad[FDLS_DEFORM_MAX_HEADS];
  size_t total_bytes;
  DmChunk c[DM_MAX_CHUNKS];
};

static int dm_npad(int k) { return ((k + 15) / 16) * 16; }

static int dm_make_plan(const FdlsDeformArgs* a, DmPlan* pl) {
  if (!a || a->G < 16 || a->G > 128 || (a->G % 16) || (a->W != 64 && a->W != 128) || a->H < 1 ||
      a->H > FDLS_DEFORM_MAX_HEADS)
    return FDLS_EUNSUPPORTED;
  pl->n_chunks = 0;
  size_t off = 0;
  auto add = [&](int kind, int head, int N, int K) {
    const int total = K / 8;
    int per = DM_CHUNK_BYTES / (N * 16 * PARTS);  // slabs per chunk
    per &= ~1;
    if (per < 2) return false;
    for (int s0 = 0; s0 < total; s0 += per) {
      const int n = total - s0 < per ? total - s0 : per;
      DmChunk& c = pl->c[pl->n_chunks++];
      c.off = (uint32_t)off;
      c.bytes = (uint32_t)(PARTS * n * N * 16);
      c.N = (uint16_t)N; c.slab0 = (uint16_t)s0; c.slabs = (uint16_t)n;
      c.kind = (uint8_t)kind; c.head = (uint8_t)head;
      c.first = s0 == 0; c.last = s0 + n >= total;
      off += align_up(c.bytes, 128);
    }
    return true;
  };
  if (!add(DM_KIND_TRUNK, 0, a->W, a->G)) return FDLS_EUNSUPPORTED;
  for (int h = 0; h < a->H; ++h) {
    const int k = a->head[h].k;
    if (k < 1 || k > 48) return FDLS_EUNSUPPORTED;
    pl->npad[h] = dm_npad(k);
    if (!add(DM_KIND_HIDDEN, h, a->W, a->W)) return FDLS_EUNSUPPORTED;
    if (!add(DM_KIND_OUT, h, pl->npad[h], a->W)) return FDLS_EUNSUPPORTED;
  }
  pl->total_bytes = off;
  return FDLS_OK;
}

// one weight matrix (N x K fp32, torch layout) -> the chunks of its GEMM: part p of chunk c at
// c.off + p * (c.slabs * Npad * 16) + (slab - c.slab0) * Npad * 16 + n * 16; rows n >= N are zero
__global__ void __launch_bounds__(256)
deform_prep_kernel(const float* __restrict__ Wm, int N, int Npad, int K, char* __restrict__ prep, DmPlan pl, int kind,
                   int head) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Npad * (K / 8)) return;
  const int n = i % Npad, slab = i / Npad;
  float v[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) v[e] = n < N ? Wm[(size_t)n * K + slab * 8 + e] : 0.f;
  for (int c = 0; c < pl.n_chunks; ++c) {
    const DmChunk& ch = pl.c[c];
    if (ch.kind == kind && ch.head == head && slab >= ch.slab0 && slab < ch.slab0 + ch.slabs) {
      store_row8_split(prep + ch.off + (size_t)(slab - ch.slab0) * Npad * 16 + n * 16, (size_t)ch.slabs * Npad * 16, v);
      return;
    }
  }
}

struct DmHeadDev {
  const float* b1; const float* b2; const float* base; float* out;
  int k, act, npad;
};
struct DmDev {
  int P, G, W, H;
  const float* feat; const float* b0; const char* prep;
  DmHeadDev head[FDLS_DEFORM_MAX_HEADS];
};

constexpr int DM_GROUP_THREADS = 256;                      // one epilogue group: two threads per point, W/2 columns each
constexpr int DM_EPI_THREADS = 2 * DM_GROUP_THREADS;       // two groups = two tiles in flight per CTA
constexpr int DM_THREADS = DM_EPI_THREADS + 32;            // + one producer warp: weight copies and MMA issue
constexpr int DM_SLOTS = 5;                                // weight-chunk ring: chunk g+3 is copied while chunk g multiplies
constexpr int DM_AHEAD = 3;
constexpr int DM_MAX_GEMMS = 1 + 2 * FDLS_DEFORM_MAX_HEADS;

template <int N>
__device__ __forceinline__ void dm_cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// like dm_issue_gemm, with the descriptors advanced by adding to their address field (3 instructions per MMA)
__device__ __forceinline__ void dm_issue_chunk(uint32_t tmem_d, uint32_t a_base, uint32_t a_part, uint32_t b_base,
                                               uint32_t b_part, int N, int ksteps, bool accumulate_first) {
  const uint32_t idesc = instr_desc_f16(DM_ROWS, N);
  const uint32_t a_slab = DM_ROWS * 16, b_slab = (uint32_t)N * 16;
  uint64_t da[PARTS], db[PARTS];
#pragma unroll
  for (int p = 0; p < PARTS; ++p) {
    da[p] = smem_desc(a_base + p * a_part, a_slab, 128);
    db[p] = smem_desc(b_base + p * b_part, b_slab, 128);
  }
  const uint64_t a_inc = (2 * a_slab) >> 4, b_inc = (2 * b_slab) >> 4;  // one K = 16 step, in 16-byte units
  bool acc = accumulate_first;
  for (int ks = 0; ks < ksteps; ++ks) {
#pragma unroll
    for (int t = 0; t < DM_NTERMS; ++t) {
      mma_f16(tmem_d, da[DM_TERMS[t][0]], db[DM_TERMS[t][1]], idesc, acc);
      acc = true;
    }
#pragma unroll
    for (int p = 0; p < PARTS; ++p) {
      da[p] += a_inc;
      db[p] += b_inc;
    }
  }
}

// The producer's walk over (pair of tiles, GEMM, tile slot, chunk): GEMM q of slot 0, then GEMM q of slot 1, then q+1 ...
// so that slot 1's MMAs run while group 0 turns slot 0's accumulator into its next operand, and vice versa.
struct DmWalk {
  int pair, q, s, ci;
  __device__ __forceinline__ bool valid(int n_pairs) const { return pair < n_pairs; }
  __device__ __forceinline__ void next(const int* gemm_count, int n_gemms, int my_tiles) {
    if (++ci < gemm_count[q]) return;
    ci = 0;
    const int slots = 2 * pair + 1 < my_tiles ? 2 : 1;
    if (++s < slots) return;
    s = 0;
    if (++q < n_gemms) return;
    q = 0;
    ++pair;
  }
};

// Warp-specialised, two tiles in flight: the PRODUCER warp streams the weight chunks (cp.async ring) and issues every
// MMA as soon as its operands exist; two EPILOGUE groups of 8 warps each own one tile slot (its activation buffer in
// shared memory, its 256 TMEM columns) and turn accumulators into the next GEMM's A operand or into outputs.  Per
// slot, two mbarriers that flip once per GEMM:
//   a_ready[s]   (256 arrivals): the A operand of the slot's next GEMM is in shared memory      group s -> producer
//   acc_ready[s] (tcgen05.commit): the accumulator is complete and the activation buffer free   producer -> group s
__global__ void __launch_bounds__(DM_THREADS, 1)
deform_mlp_kernel(DmDev a, DmPlan pl) {
  extern __shared__ __align__(128) unsigned char dm_smem[];
  __shared__ uint64_t s_done[DM_SLOTS];  // MMAs of the chunk in ring slot i have completed
  __shared__ uint64_t s_full[DM_SLOTS];  // the chunk's bytes have arrived in ring slot i (TMA transaction count)
  __shared__ uint64_t s_acc_ready[2], s_a_ready[2];
  __shared__ uint32_t s_tmem;
  __shared__ float s_bias[(1 + FDLS_DEFORM_MAX_HEADS) * 128];  // b0, then b1 of every head
  __shared__ int s_gemm_first[DM_MAX_GEMMS], s_gemm_count[DM_MAX_GEMMS];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int W = a.W, G = a.G;
  const uint32_t x_part = (uint32_t)(G / 8) * DM_ROWS * 16, a_part = (uint32_t)(W / 8) * DM_ROWS * 16;
  char* sW = reinterpret_cast<char*>(dm_smem) + 2 * PARTS * a_part;  // behind the two activation buffers
  const int n_gemms = 1 + 2 * a.H;

  if (warp == 0) tmem_alloc<512>(&s_tmem);
  if (tid == 0) {
    for (int i = 0; i < DM_SLOTS; ++i) {
      mbar_init(&s_done[i], 1);
      mbar_init(&s_full[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_acc_ready[i], 1);
      mbar_init(&s_a_ready[i], DM_GROUP_THREADS);
    }
    fence_mbar_init();
    int q = -1;
    for (int c = 0; c < pl.n_chunks; ++c) {  // chunks of one GEMM are consecutive in the plan
      if (pl.c[c].first) {
        ++q;
        s_gemm_first[q] = c;
        s_gemm_count[q] = 0;
      }
      ++s_gemm_count[q];
    }
  }
  for (int i = tid; i < W; i += DM_THREADS) s_bias[i] = a.b0[i];
  for (int h = 0; h < a.H; ++h)
    for (int i = tid; i < W; i += DM_THREADS) s_bias[(1 + h) * 128 + i] = a.head[h].b1[i];
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = s_tmem;

  const int n_tiles = (a.P + DM_ROWS - 1) / DM_ROWS;
  const int my_tiles = blockIdx.x < n_tiles ? (n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const int n_pairs = (my_tiles + 1) / 2;

  if (warp == DM_EPI_THREADS / 32) {
    // =================================================== PRODUCER ===================================================
    // One thread: a chunk is ONE TMA bulk copy (cp.async.bulk, completion in bytes on s_full[slot]) and 3 x ksteps MMAs.
    // (A cp.async loop over the warp's lanes cost ~300 instructions per chunk and made this warp, not the tensor
    // pipe, the bound of the kernel.)
    if (lane == 0) {
      DmWalk pf = {0, 0, 0, 0}, cur = {0, 0, 0, 0};
      long long pf_step = 0;
      auto prefetch = [&]() {  // copy the chunk `pf` points at into ring slot pf_step % SLOTS
        if (pf.valid(n_pairs)) {
          const DmChunk& ch = pl.c[s_gemm_first[pf.q] + pf.ci];
          uint64_t* bar = &s_full[pf_step % DM_SLOTS];
          mbar_expect_tx(bar, ch.bytes);
          bulk_copy_g2s(sW + (size_t)(pf_step % DM_SLOTS) * DM_CHUNK_BYTES, a.prep + ch.off, ch.bytes, bar);
          pf.next(s_gemm_count, n_gemms, my_tiles);
        }
        ++pf_step;
      };
      for (int i = 0; i < DM_AHEAD; ++i) prefetch();
      uint32_t gemm_no[2] = {0, 0};  // GEMMs started per slot
      for (long long g = 0; cur.valid(n_pairs); ++g) {
        const DmChunk& ch = pl.c[s_gemm_first[cur.q] + cur.ci];
        const int s = cur.s;
        // refill the ring: slot (g + AHEAD) % SLOTS held chunk g + AHEAD - SLOTS, whose MMAs must have completed
        if (g >= DM_SLOTS - DM_AHEAD)
          mbar_wait(&s_done[(g + DM_AHEAD) % DM_SLOTS], (uint32_t)(((g + DM_AHEAD - DM_SLOTS) / DM_SLOTS) & 1));
        prefetch();
        mbar_wait(&s_full[g % DM_SLOTS], (uint32_t)((g / DM_SLOTS) & 1));  // chunk g has landed
        if (ch.first) {  // this GEMM's A operand: written by group s (fenced by its threads before they arrived)
          mbar_wait(&s_a_ready[s], gemm_no[s] & 1);
          ++gemm_no[s];
        }
        tc_fence_after_sync();
        const uint32_t a_base = smem_u32(dm_smem) + (uint32_t)s * PARTS * a_part;  // trunk: the X tile aliases it
        const uint32_t a_prt = ch.kind == DM_KIND_TRUNK ? x_part : a_part;
        const uint32_t d_col = (uint32_t)s * 256u + (ch.kind == DM_KIND_TRUNK ? 0u : (uint32_t)W);
        const uint32_t b_base = smem_u32(sW + (size_t)(g % DM_SLOTS) * DM_CHUNK_BYTES);
        dm_issue_chunk(tmem + d_col, a_base + (uint32_t)ch.slab0 * DM_ROWS * 16, a_prt, b_base,
                       (uint32_t)ch.slabs * ch.N * 16, ch.N, ch.slabs / 2, !ch.first);
        mma_commit(&s_done[g % DM_SLOTS]);
        if (ch.last) mma_commit(&s_acc_ready[s]);
        cur.next(s_gemm_count, n_gemms, my_tiles);
      }
    }
  } else {
    // =================================================== EPILOGUE ===================================================
    const int grp = tid / DM_GROUP_THREADS;  // tile slot of this group
    const int eid = tid % DM_GROUP_THREADS;
    const int prow = eid & (DM_ROWS - 1);    // point of the tile = TMEM lane
    const int half = eid >> 7;               // which half of the columns / slabs this thread takes
    char* sA = reinterpret_cast<char*>(dm_smem) + (size_t)grp * PARTS * a_part;
    char* sX = sA;  // the input tile is dead once the trunk GEMM has completed, which is when sA is first written
    const uint32_t tbase = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)grp * 256u;  // lane quadrant warp % 4
    uint32_t gemm = 0;  // accumulators consumed so far
    auto operand_ready = [&]() {  // my part of the next GEMM's A operand is written
      fence_proxy_async();
      mbar_arrive(&s_a_ready[grp]);
    };
    auto wait_acc = [&]() {
      mbar_wait(&s_acc_ready[grp], gemm & 1);
      ++gemm;
      tc_fence_after_sync();
    };
    // TMEM columns [col, col + W) of the slot + bias, ReLU -> sA; this thread: W/2 columns of its point
    auto relu_to_A = [&](uint32_t col, const float* __restrict__ bias) {
      const int c0 = half * (W / 2);
      for (int c = c0; c < c0 + W / 2; c += 32) {
        float v[32];
        tmem_ld32(tbase + col + c, v);
#pragma unroll
        for (int sl = 0; sl < 4; ++sl) {
          float w8[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) w8[e] = fmaxf(v[sl * 8 + e] + bias[c + sl * 8 + e], 0.f);
          store_row8_split(sA + (size_t)(c / 8 + sl) * DM_ROWS * 16 + prow * 16, a_part, w8);
        }
      }
      tc_fence_before_sync();
    };
    auto write_out = [&](int h, int row) {  // the group's first four warps (half 0); the outputs alias the hidden columns
      const DmHeadDev& hd = a.head[h];
      float y[48];
      for (int c = 0; c < hd.npad; c += 8) {
        float v[8];
        tmem_ld8(tbase + W + c, v);
#pragma unroll
        for (int e = 0; e < 8; ++e) y[c + e] = v[e];
      }
      tc_fence_before_sync();
      if (row >= a.P) return;
      float nrm = 0.f;
      for (int e = 0; e < hd.k; ++e) {
        float t = y[e] + __ldg(hd.b2 + e);
        if (hd.base) t = hd.base[(size_t)row * hd.k + e] + t;
        if (hd.act == FDLS_DEFORM_ACT_EXP) t = expf(t);
        else if (hd.act == FDLS_DEFORM_ACT_SIGMOID) t = 1.f / (1.f + expf(-t));
        y[e] = t;
        nrm += t * t;
      }
      if (hd.act == FDLS_DEFORM_ACT_NORMALIZE) {
        const float inv = 1.f / fmaxf(sqrtf(nrm), 1e-12f);
        for (int e = 0; e < hd.k; ++e) y[e] *= inv;
      }
      for (int e = 0; e < hd.k; ++e) hd.out[(size_t)row * hd.k + e] = y[e];
    };

    for (int t = grp; t < my_tiles; t += 2) {
      const int row = ((int)blockIdx.x + t * (int)gridDim.x) * DM_ROWS + prow;
      // ---- X tile (aliases sA): every MMA that read sA for the previous tile has completed (last wait_acc below)
      for (int slab = half; slab < G / 8; slab += 2) {
        float v[8];
        if (row < a.P) {
          const float4 p = *reinterpret_cast<const float4*>(a.feat + (size_t)row * G + slab * 8);
          const float4 q = *reinterpret_cast<const float4*>(a.feat + (size_t)row * G + slab * 8 + 4);
          v[0] = p.x; v[1] = p.y; v[2] = p.z; v[3] = p.w; v[4] = q.x; v[5] = q.y; v[6] = q.z; v[7] = q.w;
        } else {
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] = 0.f;
        }
        store_row8_split(sX + (size_t)slab * DM_ROWS * 16 + prow * 16, x_part, v);
      }
      operand_ready();  // trunk GEMM may start
      for (int h = 0; h < a.H; ++h) {
        wait_acc();  // trunk (h = 0) or the previous head's layer 2: its accumulator is complete and sA is free
        if (h > 0 && half == 0) write_out(h - 1, row);  // before layer 1 of head h overwrites these columns
        relu_to_A(0, s_bias);
        operand_ready();  // layer 1 of head h
        wait_acc();       // layer 1
        relu_to_A((uint32_t)W, s_bias + (1 + h) * 128);
        operand_ready();  // layer 2 of head h (its output aliases the hidden columns just consumed)
      }
      wait_acc();  // layer 2 of the last head
      if (half == 0) write_out(a.H - 1, row);
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tmem);
}

static size_t dm_smem_bytes(int G, int W) {
  const int K = G > W ? G : W;  // the input tile aliases the activation buffer
  return (size_t)2 * PARTS * (K / 8) * DM_ROWS * 16 + (size_t)DM_SLOTS * DM_CHUNK_BYTES;  // two tile slots
}

size_t deform_prep_bytes(const FdlsDeformArgs* a) {
  DmPlan pl;
  return dm_make_plan(a, &pl) == FDLS_OK ? pl.total_bytes : 0;
}

int deform_prepare(const FdlsDeformArgs* a, cudaStream_t st) {
  DmPlan pl;
  int rc = dm_make_plan(a, &pl);
  if (rc) return rc;
  if (!a->prep || a->prep_bytes < pl.total_bytes || !a->w0) return FDLS_EINVAL;
  char* prep = (char*)a->prep;
  auto run = [&](const float* Wm, int N, int Npad, int K, int kind, int head) {
    const int n = Npad * (K / 8);
    deform_prep_kernel<<<cdiv(n, 256), 256, 0, st>>>(Wm, N, Npad, K, prep, pl, kind, head);
    count_launch();
  };
  run(a->w0, a->W, a->W, a->G, DM_KIND_TRUNK, 0);
  for (int h = 0; h < a->H; ++h) {
    if (!a->head[h].w1 || !a->head[h].w2) return FDLS_EINVAL;
    run(a->head[h].w1, a->W, a->W, a->W, DM_KIND_HIDDEN, h);
    run(a->head[h].w2, a->head[h].k, pl.npad[h], a->W, DM_KIND_OUT, h);
  }
  return cudaGetLastError() == cudaSuccess ? FDLS_OK : FDLS_ECUDA;
}

int deform_forward(const FdlsDeformArgs* a, cudaStream_t st) {
  DmPlan pl;
  int rc = dm_make_plan(a, &pl);
  if (rc) return rc;
  if (a->P < 0 || !a->prep || a->prep_bytes < pl.total_bytes) return FDLS_EINVAL;
  if (a->P == 0) return FDLS_OK;
  if (!a->feat || !a->b0) return FDLS_EINVAL;
  DmDev d;
  d.P = a->P; d.G = a->G; d.W = a->W; d.H = a->H;
  d.feat = a->feat; d.b0 = a->b0; d.prep = (const char*)a->prep;
  for (int h = 0; h < a->H; ++h) {
    const FdlsDeformHead& s = a->head[h];
    if (!s.b1 || !s.b2 || !s.out) return FDLS_EINVAL;
    d.head[h] = {s.b1, s.b2, s.base, s.out, s.k, s.act, pl.npad[h]};
  }
  const size_t smem = dm_smem_bytes(a->G, a->W);
  static SmemOptIn optin;
  if (optin.ensure(deform_mlp_kernel, smem) != cudaSuccess) return FDLS_ECUDA;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int n_tiles = cdiv(a->P, DM_ROWS);
  count_launch();
  deform_mlp_kernel<<<n_tiles < sms ? n_tiles : sms, DM_THREADS, smem, st>>>(d, pl);
  return cudaGetLastError() == cudaSuccess ? FDLS_OK : FDLS_ECUDA;
}

}  // namespace fdls
