// train_ops.cu -- the small HBM-bound kernels either side of the rasterizer in a training iteration:
//   * masked L1 loss, forward and image gradient in ONE pass              (train.py:283-296, utils/loss_utils.py:20-21)
//   * ground-truth language-feature map: segment-id -> feature gather     (scene/cameras.py:97-118)
//   * Adam over many tensors in one launch                                (scene/gaussian_model.py:301, torch.optim.Adam)
//   * densification statistics and row compaction                         (scene/gaussian_model.py:746-748, :487-520)
// Every kernel is a single streaming pass: 16-byte accesses where the layout allows, grid sized to cover the array once.
#include "common.cuh"

namespace fdls {

// ---- masked L1 ---------------------------------------------------------------------------------------------------------
// loss += weight * sum |pred*m - gt*m| ; dL_dpred = weight * sign(pred*m - gt*m) * m     (m = 1 when mask is null)
// pred, gt: (C, HW) planar; mask: (HW) floats shared by the channels (language_feature_mask, (1,H,W) in the reference).
// `weight` carries the mean's 1/N and any loss coefficient (args.lam), so the gradient is final: the rasterizer's
// backward reads it as dL_dout_* without another elementwise pass.
__global__ void __launch_bounds__(256)
masked_l1_kernel(int C, long long HW, const float* __restrict__ pred, const float* __restrict__ gt,
                 const float* __restrict__ mask, float weight, float* __restrict__ grad, float* __restrict__ loss) {
  pdl_grid_sync();
  __shared__ float s_part[8];
  const long long n4 = HW / 4;
  float acc = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 m = mask ? reinterpret_cast<const float4*>(mask)[i] : make_float4(1.f, 1.f, 1.f, 1.f);
    for (int c = 0; c < C; ++c) {
      const float4 p = reinterpret_cast<const float4*>(pred + (size_t)c * HW)[i];
      const float4 g = reinterpret_cast<const float4*>(gt + (size_t)c * HW)[i];
      const float d0 = p.x * m.x - g.x * m.x, d1 = p.y * m.y - g.y * m.y, d2 = p.z * m.z - g.z * m.z,
                  d3 = p.w * m.w - g.w * m.w;
      acc += fabsf(d0) + fabsf(d1) + fabsf(d2) + fabsf(d3);
      if (grad) {
        auto sg = [](float d) { return d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f); };  // torch's sign(0) = 0
        reinterpret_cast<float4*>(grad + (size_t)c * HW)[i] =
            make_float4(weight * sg(d0) * m.x, weight * sg(d1) * m.y, weight * sg(d2) * m.z, weight * sg(d3) * m.w);
      }
    }
  }
  // tail (HW % 4) by the first threads of block 0
  if (blockIdx.x == 0 && threadIdx.x < (int)(HW - n4 * 4)) {
    const long long i = n4 * 4 + threadIdx.x;
    const float m = mask ? mask[i] : 1.f;
    for (int c = 0; c < C; ++c) {
      const float d = pred[(size_t)c * HW + i] * m - gt[(size_t)c * HW + i] * m;
      acc += fabsf(d);
      if (grad) grad[(size_t)c * HW + i] = weight * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f)) * m;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < 8 ? s_part[threadIdx.x] : 0.f;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) atomicAdd(loss, weight * v);
  }
}

// any size / alignment: one element per thread step
__global__ void __launch_bounds__(256)
masked_l1_scalar_kernel(int C, long long HW, const float* __restrict__ pred, const float* __restrict__ gt,
                        const float* __restrict__ mask, float weight, float* __restrict__ grad, float* __restrict__ loss) {
  pdl_grid_sync();
  float acc = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < HW; i += (long long)gridDim.x * blockDim.x) {
    const float m = mask ? mask[i] : 1.f;
    for (int c = 0; c < C; ++c) {
      const float d = pred[(size_t)c * HW + i] * m - gt[(size_t)c * HW + i] * m;
      acc += fabsf(d);
      if (grad) grad[(size_t)c * HW + i] = weight * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f)) * m;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) atomicAdd(loss, weight * acc);
}

int masked_l1(int C, long long HW, const float* pred, const float* gt, const float* mask, float weight, float* grad,
              float* loss, cudaStream_t st) {
  if (C < 1 || HW < 0 || !pred || !gt || !loss) return FDLS_EINVAL;
  if (HW == 0) return FDLS_OK;
  // 16-byte accesses need every plane 16-byte aligned (HW % 4 == 0 and aligned bases); otherwise the scalar kernel
  const bool vec = (HW % 4 == 0) && (((uintptr_t)pred | (uintptr_t)gt | (uintptr_t)mask | (uintptr_t)grad) % 16 == 0);
  const long long n4 = vec ? HW / 4 : HW;
  int grid = (int)((n4 + 255) / 256);
  if (grid > 148 * 8) grid = 148 * 8;
  count_launch();
  if (vec) launch_pdl(masked_l1_kernel, dim3(grid), dim3(256), 0, st, C, HW, pred, gt, mask, weight, grad, loss);
  else launch_pdl(masked_l1_scalar_kernel, dim3(grid), dim3(256), 0, st, C, HW, pred, gt, mask, weight, grad, loss);
  return cudaGetLastError() == cudaSuccess ? FDLS_OK : FDLS_ECUDA;
}

// ---- GT language-feature gather ------------------------------------------------------------------------------------
// out[c][pix] = feature_map[seg[pix]][c], mask[pix] = seg[pix] != -1; a negative id indexes from the end like the
// reference's tensor indexing does (feature_map[-1] is the LAST row: scene/cameras.py:103 with seg = -1).
__global__ void __launch_bounds__(256)
gt_gather_kernel(long long HW, int K, int F, const int32_t* __restrict__ seg, const float* __restrict__ fmap,
                 float* __restrict__ out, uint8_t* __restrict__ mask) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= HW) return;
  const int s = seg[i];
  int row = s < 0 ? s + K : s;
  row = row < 0 ? 0 : (row >= K ? K - 1 : row);  // ids outside [-K, K) would fault in the reference; clamp instead
  for (int c = 0; c < F; ++c) out[(size_t)c * HW + i] = fmap[(size_t)row * F + c];
  if (mask) mask[i] = s != -1;
}

int gt_gather(long long HW, int K, int F, const int32_t* seg, const float* fmap, float* out, uint8_t* mask,
              cudaStream_t st) {
  if (HW < 0 || K < 1 || F < 1 || !seg || !fmap || !out) return FDLS_EINVAL;
  if (HW == 0) return FDLS_OK;
  count_launch();
  gt_gather_kernel<<<(unsigned)((HW + 255) / 256), 256, 0, st>>>(HW, K, F, seg, fmap, out, mask);
  return cudaGetLastError() == cudaSuccess ? FDLS_OK : FDLS_ECUDA;
}

// ---- Adam, many tensors, one launch -----------------------------------------------------------------------------------
// torch.optim.Adam (no weight decay, no amsgrad), the single-tensor formulation:
//   m = m + (1-b1) (g - m) ; v = b2 v + (1-b2) g g ; p -= (lr / (1-b1^t)) * m / (sqrt(v) / sqrt(1-b2^t) + eps)
constexpr int ADAM_MAX = 32;
struct AdamTable {
  float* p[ADAM_MAX]; const float* g[ADAM_MAX]; float* m[ADAM_MAX]; float* v[ADAM_MAX];
  long long n[ADAM_MAX], blk0[ADAM_MAX + 1];
  float step_size[ADAM_MAX], bc2_sqrt[ADAM_MAX], eps[ADAM_MAX];
  float b2, one_minus_b1, one_minus_b2;  // the complements are rounded from double, like torch's Python scalars
  int count;
};
__global__ void __launch_bounds__(256)
adam_kernel(AdamTable t) {
  int k = 0;
  while (k + 1 < t.count && (long long)blockIdx.x >= t.blk0[k + 1]) ++k;
  const long long i = ((long long)blockIdx.x - t.blk0[k]) * 256 + threadIdx.x;
  if (i >= t.n[k]) return;
  const float g = t.g[k][i];
  const float m0 = t.m[k][i];
  const float m = m0 + t.one_minus_b1 * (g - m0);  // exp_avg.lerp_(grad, 1 - beta1), torch's own rounding
  const float v = t.b2 * t.v[k][i] + t.one_minus_b2 * (g * g);
  t.m[k][i] = m;
  t.v[k][i] = v;
  t.p[k][i] -= t.step_size[k] * (m / (sqrtf(v) / t.bc2_sqrt[k] + t.eps[k]));
}

int adam_step(int count, float* const* p, const float* const* g, float* const* m, float* const* v, const long long* n,
              const float* lr, const float* eps, const int* step, double b1, double b2, cudaStream_t st) {
  if (count < 0 || count > ADAM_MAX) return FDLS_EINVAL;
  AdamTable t;
  t.count = 0;
  t.b2 = (float)b2;
  t.one_minus_b1 = (float)(1.0 - b1);
  t.one_minus_b2 = (float)(1.0 - b2);
  long long blocks = 0;
  for (int k = 0; k < count; ++k) {
    if (n[k] <= 0) continue;
    if (!p[k] || !g[k] || !m[k] || !v[k] || step[k] < 1) return FDLS_EINVAL;
    const int j = t.count++;
    t.p[j] = p[k]; t.g[j] = g[k]; t.m[j] = m[k]; t.v[j] = v[k]; t.n[j] = n[k];
    t.blk0[j] = blocks;
    blocks += (n[k] + 255) / 256;
    const double bc1 = 1.0 - pow(b1, step[k]), bc2 = 1.0 - pow(b2, step[k]);
    t.step_size[j] = (float)(lr[k] / bc1);
    t.bc2_sqrt[j] = (float)sqrt(bc2);
    t.eps[j] = eps[k];
  }
  t.blk0[t.count] = blocks;
  if (blocks == 0) return FDLS_OK;
  count_launch();
  adam_kernel<<<(unsigned)blocks, 256, 0, st>>>(t);
  return cudaGetLastError() == cudaSuccess ? FDLS_OK : FDLS_ECUDA;
}

// ---- densification ------------------------------------------------------------------------------------------------
// add_densification_stats (gaussian_model.py:746-748): for Gaussians in `filter`:
//   xyz_gradient_accum += ||viewspace_grad[:, :2]|| ; denom += 1
__global__ void __launch_bounds__(256)
densify_stats_kernel(int P, const float* __restrict__ vgrad, const uint8_t* __restrict__ filter,
                     float* __restrict__ accum, float* __restrict__ denom) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P || !filter[i]) return;
  const float x = vgrad[3 * i], y = vgrad[3 * i + 1];
  accum[i] += sqrtf(x * x + y * y);
  denom[i] += 1.f;
}

int densify_stats(int P, const float* vgrad, const uint8_t* filter, float* accum, float* denom, cudaStream_t st) {
  if (P < 0 || (P > 0 && (!vgrad || !filter || !accum || !denom))) return FDLS_EINVAL;
  if (P == 0) return FDLS_OK;
  count_launch();
  densify_stats_kernel<<<cdiv(P, 256), 256, 0, st>>>(P, vgrad, filter, accum, denom);
  return cudaGetLastError() == cudaSuccess ? FDLS_OK : FDLS_ECUDA;
}

// Row compaction (prune_points / _prune_optimizer, gaussian_model.py:487-520: tensor[mask] for every parameter and
// both Adam moments): rows with keep[i] != 0 move to position dest[i] (an exclusive scan of keep, computed by the
// caller once and shared by every array); `width` floats per row.
__global__ void __launch_bounds__(256)
compact_rows_kernel(long long n, int width, const uint8_t* __restrict__ keep, const int32_t* __restrict__ dest,
                    const float* __restrict__ in, float* __restrict__ out) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n * width) return;
  const long long r = e / width;
  if (keep[r]) out[(long long)dest[r] * width + (e - r * width)] = in[e];
}

int compact_rows(long long n, int width, const uint8_t* keep, const int32_t* dest, const float* in, float* out,
                 cudaStream_t st) {
  if (n < 0 || width < 1 || (n > 0 && (!keep || !dest || !in || !out))) return FDLS_EINVAL;
  if (n == 0) return FDLS_OK;
  count_launch();
  compact_rows_kernel<<<(unsigned)((n * width + 255) / 256), 256, 0, st>>>(n, width, keep, dest, in, out);
  return cudaGetLastError() == cudaSuccess ? FDLS_OK : FDLS_ECUDA;
}

}  // namespace fdls
