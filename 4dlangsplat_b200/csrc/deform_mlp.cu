// deform_mlp.cu -- the MLP half of the deformation network (scene/deformation.py:45-182) on the 5th-generation tensor
// cores: tcgen05.mma (kind::f16, bf16 x 3 operand split => fp32-grade products, see umma.cuh), accumulators in TMEM,
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

// the six bf16 x bf16 terms of an fp32-grade product, smallest first: (part of A, part of B)
__device__ constexpr int DM_TERMS[6][2] = {{2, 0}, {0, 2}, {1, 1}, {1, 0}, {0, 1}, {0, 0}};

// Issue D[128 x N] (+)= A[128 x K] * B[N x K]^T over `ksteps` K=16 steps.  A, B: three parts each, `a_part` / `b_part`
// bytes apart, canonical slabs (umma.cuh).  One thread.
__device__ __forceinline__ void dm_issue_gemm(uint32_t tmem_d, uint32_t a_base, uint32_t a_part, uint32_t b_base,
                                              uint32_t b_part, int N, int ksteps, bool accumulate_first) {
  const uint32_t idesc = instr_desc_bf16(DM_ROWS, N);
  const uint32_t a_slab = DM_ROWS * 16, b_slab = (uint32_t)N * 16;
  bool acc = accumulate_first;
  for (int ks = 0; ks < ksteps; ++ks) {
#pragma unroll
    for (int t = 0; t < 6; ++t) {
      const uint64_t da = smem_desc(a_base + DM_TERMS[t][0] * a_part + ks * 2 * a_slab, a_slab, 128);
      const uint64_t db = smem_desc(b_base + DM_TERMS[t][1] * b_part + ks * 2 * b_slab, b_slab, 128);
      mma_bf16(tmem_d, da, db, idesc, acc);
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
  const uint32_t a_part = (uint32_t)(K / 8) * DM_ROWS * 16;  // bytes of one bf16 part of the activation tile
  const uint32_t b_part = (uint32_t)(K / 8) * N * 16;
  char* sA = reinterpret_cast<char*>(dm_smem);
  char* sB = sA + 3 * a_part;

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
    store_row8_bf3(sB + (size_t)slab * N * 16 + n * 16, b_part, v);
  }
  // activations: thread = row
  const int row = blockIdx.x * DM_ROWS + tid;
  for (int slab = 0; slab < K / 8; ++slab) {
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = row < rows ? X[(size_t)row * K + slab * 8 + e] : 0.f;
    store_row8_bf3(sA + (size_t)slab * DM_ROWS * 16 + tid * 16, a_part, v);
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
  const size_t smem = (size_t)3 * (K / 8) * 16 * (DM_ROWS + N);
  static SmemOptIn optin;
  if (optin.ensure(umma_linear_kernel, smem) != cudaSuccess) return FDLS_ECUDA;
  count_launch();
  umma_linear_kernel<<<cdiv(rows, DM_ROWS), DM_ROWS, smem, st>>>(rows, N, K, X, W, bias, Y, relu);
  return cudaGetLastError() == cudaSuccess ? FDLS_OK : FDLS_ECUDA;
}

}  // namespace fdls
