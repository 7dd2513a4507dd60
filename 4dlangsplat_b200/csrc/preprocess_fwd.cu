// preprocess_fwd.cu -- per-Gaussian forward stage: near cull, projection, 3D covariance, EWA 2D covariance, conic,
// screen radius, tile rectangle, SH -> RGB, and packing of the 64-byte render record.
//
// Replaces preprocessCUDA<3> (cuda_rasterizer/forward.cu:155-256) with its helpers in_frustum (auxiliary.h:139-164),
// computeCov3D (forward.cu:118-152), computeCov2D (forward.cu:74-113), computeColorFromSH (forward.cu:20-71).
// Expression order follows the reference so that depth bits, radii and tile rectangles are bit-identical.
//
// B200 notes: one thread per Gaussian, 256-thread blocks; the two 4x4 matrices are staged in shared memory once per
// block; SH coefficients are fetched as 12 x 16-byte loads per Gaussian; everything the compositor needs is written
// as ONE 64-byte record (4 x st.v4) so the render kernels gather a Gaussian with two 32-byte sectors.
#include <cstdio>
#include "common.cuh"

namespace fdls {

__device__ __forceinline__ V3 sh_to_rgb(int deg, const float* __restrict__ sh, V3 dir, uint8_t& clamp_bits) {
  // forward.cu:20-71.  `sh` points at this Gaussian's (M,3) block.
  auto SH = [&](int i) { return V3{sh[3 * i], sh[3 * i + 1], sh[3 * i + 2]}; };
  V3 result = SH_C0 * SH(0);
  if (deg > 0) {
    float x = dir.x, y = dir.y, z = dir.z;
    result = result - SH_C1 * y * SH(1) + SH_C1 * z * SH(2) - SH_C1 * x * SH(3);
    if (deg > 1) {
      float xx = x * x, yy = y * y, zz = z * z;
      float xy = x * y, yz = y * z, xz = x * z;
      result = result + SH_C2[0] * xy * SH(4) + SH_C2[1] * yz * SH(5) + SH_C2[2] * (2.0f * zz - xx - yy) * SH(6) +
               SH_C2[3] * xz * SH(7) + SH_C2[4] * (xx - yy) * SH(8);
      if (deg > 2) {
        result = result + SH_C3[0] * y * (3.0f * xx - yy) * SH(9) + SH_C3[1] * xy * z * SH(10) +
                 SH_C3[2] * y * (4.0f * zz - xx - yy) * SH(11) +
                 SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * SH(12) +
                 SH_C3[4] * x * (4.0f * zz - xx - yy) * SH(13) + SH_C3[5] * z * (xx - yy) * SH(14) +
                 SH_C3[6] * x * (xx - 3.0f * yy) * SH(15);
      }
    }
  }
  result.x += 0.5f;
  result.y += 0.5f;
  result.z += 0.5f;
  clamp_bits = (result.x < 0 ? 1 : 0) | (result.y < 0 ? 2 : 0) | (result.z < 0 ? 4 : 0);
  return {fmaxf(result.x, 0.0f), fmaxf(result.y, 0.0f), fmaxf(result.z, 0.0f)};
}

// forward.cu:118-152 (quaternion deliberately NOT normalised, :127)
__device__ __forceinline__ void cov3d_from_scale_rot(float3 s, float mod, float4 q, float* cov3D) {
  M3 S = {{{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}};
  S.m[0][0] = mod * s.x;
  S.m[1][1] = mod * s.y;
  S.m[2][2] = mod * s.z;
  float r = q.x, x = q.y, y = q.z, z = q.w;
  M3 R = {{{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
           {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
           {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}}};
  M3 Mm = mul(S, R);
  M3 Sigma = mul(transpose(Mm), Mm);
  cov3D[0] = Sigma.m[0][0];
  cov3D[1] = Sigma.m[0][1];
  cov3D[2] = Sigma.m[0][2];
  cov3D[3] = Sigma.m[1][1];
  cov3D[4] = Sigma.m[1][2];
  cov3D[5] = Sigma.m[2][2];
}

// forward.cu:74-113
__device__ __forceinline__ float3 cov2d_ewa(const float3& mean, float focal_x, float focal_y, float tan_fovx,
                                            float tan_fovy, const float* cov3D, const float* vm) {
  float3 t = xform4x3(mean, vm);
  const float limx = 1.3f * tan_fovx;
  const float limy = 1.3f * tan_fovy;
  const float txtz = t.x / t.z;
  const float tytz = t.y / t.z;
  t.x = min(limx, max(-limx, txtz)) * t.z;
  t.y = min(limy, max(-limy, tytz)) * t.z;
  M3 J = {{{focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z)},
           {0.0f, focal_y / t.z, -(focal_y * t.y) / (t.z * t.z)},
           {0, 0, 0}}};
  M3 Wm = {{{vm[0], vm[4], vm[8]}, {vm[1], vm[5], vm[9]}, {vm[2], vm[6], vm[10]}}};
  M3 T = mul(Wm, J);
  M3 Vrk = {{{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}}};
  M3 cov = mul(mul(transpose(T), transpose(Vrk)), T);
  cov.m[0][0] += 0.3f;
  cov.m[1][1] += 0.3f;
  return {cov.m[0][0], cov.m[0][1], cov.m[1][1]};
}

__global__ void __launch_bounds__(256)
preprocess_fwd_kernel(int P, int D, int M, int F, const float* __restrict__ means3D, const float* __restrict__ scales,
                      const float* __restrict__ rotations, const float* __restrict__ opacities,
                      const float* __restrict__ shs, const float* __restrict__ cov3D_precomp,
                      const float* __restrict__ colors_precomp, const float* __restrict__ lang, ViewParams vp,
                      int* __restrict__ radii, float4* __restrict__ rec, uint32_t* __restrict__ tiles_touched,
                      uint8_t* __restrict__ clamped, uint64_t* __restrict__ depth_keys,
                      uint32_t* __restrict__ depth_vals, int prefiltered, uint32_t sh_magic,
                      uint32_t* __restrict__ depth_hist, uint32_t* __restrict__ tickets) {
  pdl_grid_sync();
  extern __shared__ float s_sh[];
  // num_rendered[2..3]: "last block done" tickets of the scan / partition kernels that follow on this stream
  if (tickets != nullptr && blockIdx.x == 0 && threadIdx.x < 2) tickets[threadIdx.x] = 0;  // [256][3M+1] when SH are staged (sh_magic != 0), else empty
  __shared__ float s_view[16], s_proj[16], s_cam[3];
  __shared__ uint32_t s_hist[4][256];  // digit histograms of the depth keys, for the depth sort that follows
  if (depth_hist != nullptr)
    for (int i = threadIdx.x; i < 1024; i += 256) (&s_hist[0][0])[i] = 0;
  if (threadIdx.x < 16) s_view[threadIdx.x] = vp.view[threadIdx.x];
  else if (threadIdx.x < 32) s_proj[threadIdx.x - 16] = vp.proj[threadIdx.x - 16];
  else if (threadIdx.x < 35) s_cam[threadIdx.x - 32] = vp.campos != nullptr ? vp.campos[threadIdx.x - 32] : 0.f;  // only SH needs it
  __syncthreads();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int sh_stride = 3 * M + 1;
  // The Gaussian's own attributes are requested BEFORE the SH staging below, so that their global-memory latency
  // overlaps the staging's instead of following it (the kernel is latency-bound: two dependent round trips per CTA,
  // two waves of CTAs).  Values only; the arithmetic on them is unchanged.
  float3 p_orig = {0.f, 0.f, 0.f}, in_scale = {0.f, 0.f, 0.f};
  float4 in_rot = {0.f, 0.f, 0.f, 0.f};
  float in_opacity = 0.f;
  float in_lang[FDLS_MAX_LANG_CHANNELS];
#pragma unroll
  for (int i = 0; i < FDLS_MAX_LANG_CHANNELS; ++i) in_lang[i] = 0.f;
  if (idx < P) {
    p_orig = make_float3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
    if (cov3D_precomp == nullptr) {
      in_scale = make_float3(scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]);
      in_rot = reinterpret_cast<const float4*>(rotations)[idx];
    }
    in_opacity = opacities[idx];
#pragma unroll
    for (int i = 0; i < FDLS_MAX_LANG_CHANNELS; ++i)
      if (lang != nullptr && i < F) in_lang[i] = lang[(size_t)idx * F + i];
  }
  if (sh_magic != 0) {
    // Coalesced staging of this warp's 32 x (M,3) SH block: the direct per-thread reads are 192-byte strided
    // (forward.cu:20-71 reads sh[idx*max_coeffs + i]) and waste most of every sector; here each warp streams its
    // contiguous 32*3M floats with 16-byte loads and re-reads them conflict-free (row stride 3M+1 words).
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g0 = blockIdx.x * 256 + warp * 32;
    if (g0 < P) {
      const int n_fl = min(32, P - g0) * 3 * M;
      const float* src = shs + (size_t)g0 * 3 * M;
      float* dst = s_sh + warp * 32 * sh_stride;
      if (n_fl == 1536) {  // full warp, M = 16 (degree 3): all twelve 16-byte loads in flight before the first store
        const float4* src4 = reinterpret_cast<const float4*>(src);
        float4 q[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) q[i] = src4[lane + 32 * i];
#pragma unroll
        for (int i = 0; i < 12; ++i) {
          // a row is twelve 16-byte groups: group j = lane + 32 i belongs to row j / 12 (one division per group, the
          // four words go to consecutive addresses)
          const uint32_t j = lane + 32 * i;
          const uint32_t g = __umulhi(j, 357913942u);  // j / 12 for j < 2^16
          float* d = dst + g * sh_stride + (j - g * 12) * 4;
          d[0] = q[i].x; d[1] = q[i].y; d[2] = q[i].z; d[3] = q[i].w;
        }
      } else
      for (int e0 = 4 * lane; e0 < n_fl; e0 += 128) {
        float v[4];
        if (e0 + 3 < n_fl) {
          const float4 q = *reinterpret_cast<const float4*>(src + e0);
          v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) v[k] = e0 + k < n_fl ? src[e0 + k] : 0.f;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint32_t e = e0 + k;
          const uint32_t g = __umulhi(e, sh_magic);  // e / (3M)
          if ((int)e < n_fl) dst[g * sh_stride + (e - g * 3 * M)] = v[k];
        }
      }
    }
    __syncwarp();
  }

  int out_radius = 0;
  uint32_t out_tiles = 0;
  uint32_t out_depth_bits = 0xffffffffu;  // culled Gaussians sort behind every visible one (depth > 0.2 => sign bit 0)
  if (idx < P) do {
    // in_frustum (auxiliary.h:139-164): only the near-plane test is live.
    const float3 p_view = xform4x3(p_orig, s_view);
    if (p_view.z <= 0.2f) {
      if (prefiltered) {
        printf("Point is filtered although prefiltered is set. This shouldn't happen!");
        __trap();
      }
      break;
    }
    const float4 p_hom = xform4x4(p_orig, s_proj);
    const float p_w = 1.0f / (p_hom.w + 0.0000001f);
    const float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};

    float cov3D[6];
    if (cov3D_precomp != nullptr) {
#pragma unroll
      for (int i = 0; i < 6; ++i) cov3D[i] = cov3D_precomp[6 * idx + i];
    } else {
      cov3d_from_scale_rot(in_scale, vp.scale_modifier, in_rot, cov3D);
    }
    const float3 cov = cov2d_ewa(p_orig, vp.focal_x, vp.focal_y, vp.tanfovx, vp.tanfovy, cov3D, s_view);

    const float det = (cov.x * cov.z - cov.y * cov.y);
    if (det == 0.0f) break;
    const float det_inv = 1.f / det;
    const float3 conic = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};

    const float mid = 0.5f * (cov.x + cov.z);
    const float lambda1 = mid + sqrt(max(0.1f, mid * mid - det));
    const float lambda2 = mid - sqrt(max(0.1f, mid * mid - det));
    const float my_radius = ceil(3.f * sqrt(max(lambda1, lambda2)));
    const float2 point_image = {ndc2pix(p_proj.x, vp.W), ndc2pix(p_proj.y, vp.H)};
    uint2 rect_min, rect_max;
    get_rect(point_image.x, point_image.y, (int)my_radius, vp.gx, vp.gy, rect_min, rect_max);
    if ((rect_max.x - rect_min.x) * (rect_max.y - rect_min.y) == 0) break;

    V3 rgb;
    uint8_t cl = 0;
    if (colors_precomp == nullptr) {
      V3 pos = {p_orig.x, p_orig.y, p_orig.z};
      V3 cam = {s_cam[0], s_cam[1], s_cam[2]};
      V3 dir = pos - cam;
      dir = dir / sqrt(dot3(dir, dir));
      rgb = sh_magic != 0 ? sh_to_rgb(D, s_sh + threadIdx.x * sh_stride, dir, cl)
                          : sh_to_rgb(D, shs + (size_t)idx * M * 3, dir, cl);
    } else {
      rgb = {colors_precomp[3 * idx], colors_precomp[3 * idx + 1], colors_precomp[3 * idx + 2]};
    }
    clamped[idx] = cl;

    const float* l = in_lang;
    float4* r = rec + (size_t)idx * 4;
    r[0] = make_float4(point_image.x, point_image.y, conic.x, conic.y);
    r[1] = make_float4(conic.z, in_opacity, p_view.z, rgb.x);
    r[2] = make_float4(rgb.y, rgb.z, l[0], l[1]);
    r[3] = make_float4(l[2], l[3], l[4], l[5]);
    out_radius = my_radius;
    out_tiles = (rect_max.y - rect_min.y) * (rect_max.x - rect_min.x);
    out_depth_bits = __float_as_uint(p_view.z);
  } while (0);
  if (idx < P) {
    radii[idx] = out_radius;
    tiles_touched[idx] = out_tiles;
    if (depth_keys != nullptr) {  // input of the P-element depth sort (api.cu)
      depth_keys[idx] = out_depth_bits;
      depth_vals[idx] = idx;
    }
  }
  if (depth_hist != nullptr) {  // replaces the sort's own histogram pass over the keys (radix_sort.cu)
    __syncthreads();
    if (idx < P) {
#pragma unroll
      for (int p = 0; p < 4; ++p) atomicAdd(&s_hist[p][(out_depth_bits >> (8 * p)) & 255u], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 1024; i += 256) {
      const uint32_t c = (&s_hist[0][0])[i];
      if (c) atomicAdd(&depth_hist[i], c);
    }
  }
}

// markVisible / checkFrustum (rasterizer_impl.cu:54-66)
__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                                    uint8_t* __restrict__ present) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const float3 p = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
  const float3 pv = xform4x3(p, view);
  present[idx] = pv.z > 0.2f ? 1 : 0;
}

cudaError_t launch_preprocess_fwd(const FdlsFwdArgs* a, const ViewParams& vp, float4* rec, uint32_t* tiles_touched,
                                  uint8_t* clamped, uint64_t* depth_keys, uint32_t* depth_vals, uint32_t* depth_hist,
                                  uint32_t* tickets, cudaStream_t st) {
  if (a->P == 0) return cudaSuccess;
  // SH staging needs 256*(3M+1) floats of shared memory; M <= 16 (degree 3) -> 50 KB
  const bool stage = a->shs != nullptr && a->colors_precomp == nullptr && a->M >= 1 && a->M <= 16;
  const size_t smem = stage ? (size_t)256 * (3 * a->M + 1) * sizeof(float) : 0;
  const uint32_t magic = stage ? (uint32_t)(0xffffffffu / (uint32_t)(3 * a->M) + 1u) : 0u;
  static SmemOptIn optin;  // per device (common.cuh)
  if (cudaError_t e = optin.ensure(preprocess_fwd_kernel, smem)) return e;
  launch_pdl(preprocess_fwd_kernel, dim3(cdiv(a->P, 256)), dim3(256), smem, st, a->P, a->D, a->M, a->F, a->means3D, a->scales, a->rotations, a->opacities, a->shs, a->cov3D_precomp, a->colors_precomp, a->language_feature, vp, a->radii, rec, tiles_touched, clamped, depth_keys, depth_vals, a->prefiltered, magic, depth_hist, tickets); count_launch();
  return cudaGetLastError();
}

cudaError_t launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t st) {
  if (P == 0) return cudaSuccess;
  mark_visible_kernel<<<cdiv(P, 256), 256, 0, st>>>(P, means3D, view, present); count_launch();
  return cudaGetLastError();
}

}  // namespace fdls
