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
  int gemm_first[1 + FDLS_DEFORM_MAX_HEADS], gemm_count[1 + FDLS_DEFORM_MAX_HEADS];  // chunks of GEMM q (consecutive)
  uint32_t w2_off[FDLS_DEFORM_MAX_HEADS];  // last-layer weights, fp32 [k rounded up to 4][W], zero rows behind k
  DmChunk c[DM_MAX_CHUNKS];
};

static int dm_npad(int k) { return ((k + 15) / 16) * 16; }
static int dm_kpad(int k) { return (k + 3) & ~3; }

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
    if (a->head[h].act == FDLS_DEFORM_ACT_NORMALIZE && k > 4) return FDLS_EUNSUPPORTED;  // normalised within one pass
    pl->npad[h] = dm_npad(k);
    if (!add(DM_KIND_HIDDEN, h, a->W, a->W)) return FDLS_EUNSUPPORTED;
  }
  for (int h = 0; h < a->H; ++h) {
    pl->w2_off[h] = (uint32_t)off;
    off += align_up((size_t)dm_kpad(a->head[h].k) * a->W * sizeof(float), 128);
  }
  pl->total_bytes = off;
  for (int c = 0, q = -1; c < pl->n_chunks; ++c) {
    if (pl->c[c].first) {
      ++q;
      pl->gemm_first[q] = c;
      pl->gemm_count[q] = 0;
    }
    ++pl->gemm_count[q];
  }
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
  const float* b1; const float* w2; const float* b2; const float* base; float* out;
  int k, act;
};
struct DmDev {
  int P, G, W, H;
  const float* feat; const float* b0; const char* prep;
  DmHeadDev head[FDLS_DEFORM_MAX_HEADS];
};

constexpr int DM_GROUP_THREADS = 256;                      // one epilogue group: two threads per point, W/2 columns each
constexpr int DM_EPI_THREADS = 2 * DM_GROUP_THREADS;       // two groups = two tiles in flight per CTA
constexpr int DM_THREADS = DM_EPI_THREADS + 64;            // + the MMA-issue warp and the TMA warp (one thread each)
constexpr int DM_SLOTS = 5;                                // weight-chunk ring
constexpr int DM_MAX_GEMMS = 1 + FDLS_DEFORM_MAX_HEADS;    // trunk + one layer-1 GEMM per head
constexpr int DM_W2_ROWS = 16;                             // last-layer rows (all heads, padded) kept in shared memory
constexpr int DM_OUT_BLOCK = 4;                            // outputs of a head's last layer handled per pass

// The walk over (pair of tiles, GEMM, tile slot, chunk) shared by the TMA thread and the MMA thread: GEMM q of slot 0,
// then GEMM q of slot 1, then q+1 ..., so that one slot's MMAs run while the other slot's group is in its epilogue.
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

__device__ __forceinline__ void dm_group_barrier(int grp) {  // the 256 threads of one epilogue group
  asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "n"(DM_GROUP_THREADS) : "memory");
}

// Warp-specialised, two 128-point tiles in flight per CTA:
//   TMA thread     streams the weight chunks of the GEMM sequence through the ring (cp.async.bulk, s_full / s_done)
//   MMA thread     issues every tcgen05.mma as soon as its operands exist and commits completions to mbarriers
//   2 groups x 8 warps   own one tile slot each: its activation buffer (shared memory) and its 256 TMEM columns
// Per tile: trunk GEMM -> cols A; ReLU(trunk + b0) -> sA ONCE, it is the A operand of EVERY head's layer 1; layer-1
// GEMMs alternate between cols B and cols A (the trunk's columns are dead once sA exists); a head's last layer
// (W -> k <= 48 outputs) runs on the CUDA cores inside the epilogue, in fp32, straight from the accumulator: as an MMA
// it cost as many instructions as layer 1 (the tensor core's time per instruction is set by the 128-row A operand,
// not by N) plus a split / store / hand-off round trip.
// mbarriers per slot: a_ready (256 arrivals: X tile or sA written), acc_ready[2] (tcgen05.commit, GEMMs alternate
// between the two so that a completion is never overtaken by the next one on the same barrier), cols_free (256
// arrivals: head h's accumulator has been consumed, head h+2 may overwrite it).
template <bool W2_SMEM>
__global__ void __launch_bounds__(DM_THREADS, 1)
deform_mlp_kernel(DmDev a, DmPlan pl) {
  extern __shared__ __align__(128) unsigned char dm_smem[];
  __shared__ uint64_t s_done[DM_SLOTS], s_full[DM_SLOTS];
  __shared__ uint64_t s_acc_ready[2][2], s_a_ready[2], s_cols_free[2];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(16) float s_bias[(1 + FDLS_DEFORM_MAX_HEADS) * 128];  // b0, then b1 of every head
  // last-layer weights of all heads (rows zero-padded to a multiple of four per head), when they fit: read as shared
  // memory broadcasts with immediate offsets instead of global loads through the L1 (a descriptor move per load)
  __shared__ __align__(16) float s_w2[W2_SMEM ? DM_W2_ROWS * 128 : 4];
  __shared__ int s_w2_row[FDLS_DEFORM_MAX_HEADS];  // first row of head h in s_w2
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int W = a.W, G = a.G;
  const uint32_t x_part = (uint32_t)(G / 8) * DM_ROWS * 16, a_part = (uint32_t)(W / 8) * DM_ROWS * 16;
  char* sW = reinterpret_cast<char*>(dm_smem) + 2 * PARTS * a_part;              // behind the two activation buffers
  float* s_part = reinterpret_cast<float*>(sW + (size_t)DM_SLOTS * DM_CHUNK_BYTES);  // [2 groups][128][DM_OUT_BLOCK]
  const int n_gemms = 1 + a.H;

  if (warp == 0) tmem_alloc<512>(&s_tmem);
  if (tid == 0) {
    for (int i = 0; i < DM_SLOTS; ++i) {
      mbar_init(&s_done[i], 1);
      mbar_init(&s_full[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_acc_ready[i][0], 1);
      mbar_init(&s_acc_ready[i][1], 1);
      mbar_init(&s_a_ready[i], DM_GROUP_THREADS);
      mbar_init(&s_cols_free[i], DM_GROUP_THREADS);
    }
    fence_mbar_init();
  }
  for (int i = tid; i < W; i += DM_THREADS) s_bias[i] = a.b0[i];
  if (W2_SMEM) {
    int r0 = 0;
    for (int h = 0; h < a.H; ++h) {
      const int rows = (a.head[h].k + 3) & ~3;
      for (int i = tid; i < rows * W; i += DM_THREADS) s_w2[r0 * W + i] = a.head[h].w2[i];
      if (tid == 0) s_w2_row[h] = r0;
      r0 += rows;
    }
  }
  for (int h = 0; h < a.H; ++h)
    for (int i = tid; i < W; i += DM_THREADS) s_bias[(1 + h) * 128 + i] = a.head[h].b1[i];
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = s_tmem;

  const int n_tiles = (a.P + DM_ROWS - 1) / DM_ROWS;
  const int my_tiles = blockIdx.x < n_tiles ? (n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const int n_pairs = (my_tiles + 1) / 2;

  if (warp == DM_EPI_THREADS / 32 + 1) {
    // ===================================================== TMA =====================================================
    // the whole warp walks the sequence (every value stays warp-uniform, i.e. in uniform registers); one elected lane
    // issues the copies
    DmWalk w = {0, 0, 0, 0};
    uint32_t ring = 0, lap = 0;
    while (w.valid(n_pairs)) {
      if (lap) mbar_wait(&s_done[ring], (lap - 1) & 1);  // the MMAs of the chunk that had this ring slot are done
      const DmChunk& ch = pl.c[pl.gemm_first[w.q] + w.ci];
      if (elect_one()) {
        mbar_expect_tx(&s_full[ring], ch.bytes);
        bulk_copy_g2s(sW + (size_t)ring * DM_CHUNK_BYTES, a.prep + ch.off, ch.bytes, &s_full[ring]);
      }
      __syncwarp();
      if (++ring == DM_SLOTS) {
        ring = 0;
        ++lap;
      }
      w.next(pl.gemm_count, n_gemms, my_tiles);
    }
  } else if (warp == DM_EPI_THREADS / 32) {
    // ===================================================== MMA =====================================================
    // Same shape: warp-uniform walk, elected issue.  (Run by a single lane, the loop cost ~300 dependent scalar
    // instructions per chunk -- register -> uniform-register moves around every descriptor -- and that one thread, not
    // the tensor core, set the kernel's time.)
    DmWalk cur = {0, 0, 0, 0};
    uint32_t ph_a = 0, ph_free = 0, ring = 0, ring_phase = 0;  // phase bits, one per tile slot
    const uint32_t act0 = smem_u32(dm_smem), ring0 = smem_u32(sW);
    while (cur.valid(n_pairs)) {
      const DmChunk& ch = pl.c[pl.gemm_first[cur.q] + cur.ci];
      const uint32_t s = (uint32_t)cur.s, q = (uint32_t)cur.q;  // q = 0: trunk; q = h + 1: layer 1 of head h
      mbar_wait(&s_full[ring], ring_phase);  // the chunk has landed
      if (ch.first) {
        if (q <= 1) {  // the X tile (trunk) or sA (first head) has been written by the slot's group
          mbar_wait(&s_a_ready[s], (ph_a >> s) & 1);
          ph_a ^= 1u << s;
        } else if (q >= 3) {  // head h = q - 1 >= 2 overwrites the accumulator head h - 2 was read from
          mbar_wait(&s_cols_free[s], (ph_free >> s) & 1);
          ph_free ^= 1u << s;
        }
      }
      tc_fence_after_sync();
      if (elect_one()) {
        // descriptors as their low words (address >> 4 | LBO >> 4 << 16); a K = 16 step advances the address field
        const uint32_t N = ch.N, a_prt = q == 0 ? x_part : a_part;  // the X tile aliases the activation buffer
        const uint32_t idesc = instr_desc_f16(DM_ROWS, (int)N);
        uint32_t da0 = ((act0 + s * PARTS * a_part + (uint32_t)ch.slab0 * DM_ROWS * 16) >> 4) | ((DM_ROWS * 16u >> 4) << 16);
        uint32_t da1 = da0 + (a_prt >> 4);
        uint32_t db0 = ((ring0 + ring * DM_CHUNK_BYTES) >> 4) | (N << 16);
        uint32_t db1 = db0 + (uint32_t)ch.slabs * N;
        const uint32_t d = tmem + s * 256u + ((q & 1) ? (uint32_t)W : 0u);  // trunk, heads 1, 3, ..: cols A
        uint32_t acc = ch.first ? 0u : 1u;
        for (int ks = ch.slabs >> 1; ks > 0; --ks) {
          mma_f16_lo(d, da1, db0, idesc, acc);  // lo * hi, hi * lo, hi * hi: small terms first
          mma_f16_lo(d, da0, db1, idesc, 1u);
          mma_f16_lo(d, da0, db0, idesc, 1u);
          acc = 1u;
          da0 += 2 * DM_ROWS * 16 >> 4;
          da1 += 2 * DM_ROWS * 16 >> 4;
          db0 += 2 * N;
          db1 += 2 * N;
        }
        mma_commit(&s_done[ring]);
        if (ch.last) mma_commit(&s_acc_ready[s][q & 1]);
      }
      __syncwarp();
      if (++ring == DM_SLOTS) {
        ring = 0;
        ring_phase ^= 1;
      }
      cur.next(pl.gemm_count, n_gemms, my_tiles);
    }
  } else {
    // =================================================== EPILOGUE ===================================================
    const int grp = tid / DM_GROUP_THREADS;  // tile slot of this group
    const int eid = tid % DM_GROUP_THREADS;
    const int prow = eid & (DM_ROWS - 1);    // point of the tile = TMEM lane
    const int half = eid >> 7;               // which half of the columns / slabs this thread takes
    char* sA = reinterpret_cast<char*>(dm_smem) + (size_t)grp * PARTS * a_part;
    char* sX = sA;  // the input tile is dead once the trunk GEMM has completed, which is when sA is first written
    float* part = s_part + (size_t)grp * DM_ROWS * DM_OUT_BLOCK;
    const uint32_t tbase = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)grp * 256u;  // lane quadrant warp % 4
    uint32_t n_acc[2] = {0, 0};  // completions consumed on each of the slot's two accumulator barriers
    auto wait_acc = [&](int q) {
      mbar_wait(&s_acc_ready[grp][q & 1], n_acc[q & 1] & 1);
      ++n_acc[q & 1];
      tc_fence_after_sync();
    };
    const int c0 = half * (W / 2);

    for (int t = grp; t < my_tiles; t += 2) {
      const int row = ((int)blockIdx.x + t * (int)gridDim.x) * DM_ROWS + prow;
      // ---- X tile (aliases sA): every MMA that read sA for the previous tile has completed (its accumulators were
      // all waited for below)
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
      fence_proxy_async();
      mbar_arrive(&s_a_ready[grp]);  // trunk GEMM may start
      // ---- trunk accumulator (cols A) + b0, ReLU -> sA: the A operand of every head's first layer
      wait_acc(0);
      for (int c = c0; c < c0 + W / 2; c += 32) {
        float v[32];
        tmem_ld32(tbase + c, v);
#pragma unroll
        for (int sl = 0; sl < 4; ++sl) {
          float w8[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) w8[e] = fmaxf(v[sl * 8 + e] + s_bias[c + sl * 8 + e], 0.f);  // LDS.128 x 2
          store_row8_split(sA + (size_t)(c / 8 + sl) * DM_ROWS * 16 + prow * 16, a_part, w8);
        }
      }
      tc_fence_before_sync();
      fence_proxy_async();
      mbar_arrive(&s_a_ready[grp]);  // layer 1 of heads 0 and 1 may start (cols B, cols A)
      // ---- per head: hidden = ReLU(acc + b1); out = act(base + hidden W2^T + b2), DM_OUT_BLOCK outputs per pass
      for (int h = 0; h < a.H; ++h) {
        const DmHeadDev& hd = a.head[h];
        const int kpad = (hd.k + 3) & ~3;  // W2 is stored with zero rows up to a multiple of 4: no guards in the FMAs
        const uint32_t col = ((h + 1) & 1) ? (uint32_t)W : 0u;
        const float4* b1 = reinterpret_cast<const float4*>(s_bias + (1 + h) * 128);
        for (int e0 = 0; e0 < kpad; e0 += DM_OUT_BLOCK) {
          // canonical attribute of this pass: requested before the accumulator wait, consumed after the FMAs
          float bs[DM_OUT_BLOCK];
#pragma unroll
          for (int e = 0; e < DM_OUT_BLOCK; ++e)
            bs[e] = (half == 0 && row < a.P && hd.base && e0 + e < hd.k) ? __ldg(hd.base + (size_t)row * hd.k + e0 + e) : 0.f;
          if (e0 == 0) wait_acc(h + 1);
          float2 acc2[DM_OUT_BLOCK];  // (sum over even columns, sum over odd columns)
#pragma unroll
          for (int e = 0; e < DM_OUT_BLOCK; ++e) acc2[e] = make_float2(0.f, 0.f);
          const float4* w2 = W2_SMEM ? reinterpret_cast<const float4*>(s_w2 + (s_w2_row[h] + e0) * W)
                                     : reinterpret_cast<const float4*>(hd.w2 + (size_t)e0 * W);
          for (int c = c0; c < c0 + W / 2; c += 16) {
            float v[16];
            tmem_ld16(tbase + col + c, v);
            float2 vp[8];
#pragma unroll
            for (int j4 = 0; j4 < 4; ++j4) {
              const float4 bv = b1[c / 4 + j4];
              vp[2 * j4] = make_float2(fmaxf(v[4 * j4] + bv.x, 0.f), fmaxf(v[4 * j4 + 1] + bv.y, 0.f));
              vp[2 * j4 + 1] = make_float2(fmaxf(v[4 * j4 + 2] + bv.z, 0.f), fmaxf(v[4 * j4 + 3] + bv.w, 0.f));
            }
#pragma unroll
            for (int ep = 0; ep < DM_OUT_BLOCK; ep += 2) {  // two outputs at a time: 8 loads in flight, then 16 FFMA2
              float4 wv[2][4];
#pragma unroll
              for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j4 = 0; j4 < 4; ++j4)  // the same address in every lane: a broadcast
                  wv[i][j4] = W2_SMEM ? w2[((ep + i) * W + c) / 4 + j4] : __ldg(w2 + ((ep + i) * W + c) / 4 + j4);
#pragma unroll
              for (int j4 = 0; j4 < 4; ++j4)
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                  acc2[ep + i] = ffma2(vp[2 * j4], make_float2(wv[i][j4].x, wv[i][j4].y), acc2[ep + i]);
                  acc2[ep + i] = ffma2(vp[2 * j4 + 1], make_float2(wv[i][j4].z, wv[i][j4].w), acc2[ep + i]);
                }
            }
          }
          float acc[DM_OUT_BLOCK];
#pragma unroll
          for (int e = 0; e < DM_OUT_BLOCK; ++e) acc[e] = acc2[e].x + acc2[e].y;
          // the two halves of a point's hidden vector live in two threads of different warps: combine through shared
          dm_group_barrier(grp);  // the previous pass's partial sums have been read
          if (half == 1) {
            *reinterpret_cast<float4*>(part + prow * DM_OUT_BLOCK) = make_float4(acc[0], acc[1], acc[2], acc[3]);
          }
          dm_group_barrier(grp);
          if (half == 0 && row < a.P) {
            const float4 p0 = *reinterpret_cast<const float4*>(part + prow * DM_OUT_BLOCK);
            const float other[DM_OUT_BLOCK] = {p0.x, p0.y, p0.z, p0.w};
            float y[DM_OUT_BLOCK];
            float nrm = 0.f;
#pragma unroll
            for (int e = 0; e < DM_OUT_BLOCK; ++e) {
              y[e] = 0.f;
              if (e0 + e < hd.k) {
                float tv = acc[e] + other[e] + __ldg(hd.b2 + e0 + e);
                if (hd.base) tv = bs[e] + tv;
                if (hd.act == FDLS_DEFORM_ACT_EXP) tv = expf(tv);
                else if (hd.act == FDLS_DEFORM_ACT_SIGMOID) tv = 1.f / (1.f + expf(-tv));
                y[e] = tv;
                nrm += tv * tv;
              }
            }
            const float inv = hd.act == FDLS_DEFORM_ACT_NORMALIZE ? 1.f / fmaxf(sqrtf(nrm), 1e-12f) : 1.f;  // k <= 4
#pragma unroll
            for (int e = 0; e < DM_OUT_BLOCK; ++e)
              if (e0 + e < hd.k) hd.out[(size_t)row * hd.k + e0 + e] = y[e] * inv;
          }
        }
        tc_fence_before_sync();
        if (h + 2 < a.H) mbar_arrive(&s_cols_free[grp]);  // head h + 2 may overwrite this accumulator
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tmem);
}

static size_t dm_smem_bytes(int G, int W) {
  const int K = G > W ? G : W;  // the input tile aliases the activation buffer
  return (size_t)2 * PARTS * (K / 8) * DM_ROWS * 16 + (size_t)DM_SLOTS * DM_CHUNK_BYTES +  // two tile slots, ring,
         (size_t)2 * DM_ROWS * DM_OUT_BLOCK * sizeof(float);                                   // partial sums
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
    const size_t row = (size_t)a->W * sizeof(float);
    if (cudaMemsetAsync(prep + pl.w2_off[h], 0, dm_kpad(a->head[h].k) * row, st) != cudaSuccess ||
        cudaMemcpyAsync(prep + pl.w2_off[h], a->head[h].w2, a->head[h].k * row, cudaMemcpyDeviceToDevice, st) != cudaSuccess)
      return FDLS_ECUDA;
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
    d.head[h] = {s.b1, (const float*)(d.prep + pl.w2_off[h]), s.b2, s.base, s.out, s.k, s.act};
  }
  const size_t smem = dm_smem_bytes(a->G, a->W);
  static SmemOptIn optin;
  int w2_rows = 0;
  for (int h = 0; h < a->H; ++h) w2_rows += dm_kpad(a->head[h].k);
  const bool w2_smem = w2_rows <= DM_W2_ROWS;
  static SmemOptIn optin_g;
  if ((w2_smem ? optin.ensure(deform_mlp_kernel<true>, smem) : optin_g.ensure(deform_mlp_kernel<false>, smem)) != cudaSuccess)
    return FDLS_ECUDA;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int n_tiles = cdiv(a->P, DM_ROWS);
  count_launch();
  if (w2_smem) deform_mlp_kernel<true><<<n_tiles < sms ? n_tiles : sms, DM_THREADS, smem, st>>>(d, pl);
  else deform_mlp_kernel<false><<<n_tiles < sms ? n_tiles : sms, DM_THREADS, smem, st>>>(d, pl);
  return cudaGetLastError() == cudaSuccess ? FDLS_OK : FDLS_ECUDA;
}

}  // namespace fdls
