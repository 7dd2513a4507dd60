// preprocess_bwd.cu -- ONE pass over the Gaussians that turns the screen-space gradient records of render_bwd into
// the API gradients.  Replaces computeCov2DCUDA (cuda_rasterizer/backward.cu:144-281) and preprocessCUDA<3> backward
// (backward.cu:353-431) with its helpers computeColorFromSH backward (:20-139) and computeCov3D backward (:285-348),
// and the eleven torch::zeros fills of RasterizeGaussiansBackwardCUDA (rasterize_points.cu:169-192): every output
// element is written exactly once here (zeros for culled Gaussians), so no memset pass precedes the kernel.
#include "common.cuh"

namespace fdls {

struct BwdOut {
  float* means3D; float* means2D; float* sh; float* colors; float* lang; float* opacity; float* scales; float* rots;
  float* cov3D; float* conic; float* depths;
  float* sh_rgb;  // (P,3) dL_dRGB after the clamp mask: the view-dependent factor of the rank-1 SH gradient (never ACC)
  int sh_rgb_header;  // also write {campos, 1.0f} into sh_rgb[-4 .. -1] (the slot header of sh_grad_expand)
};

// forward.cu:118-152 recomputed (the forward does not keep cov3D; it is 30 flops)
__device__ __forceinline__ void cov3d_build(float3 s, float mod, float4 q, float* cov3D, M3& Mm, M3& R, float3& sc) {
  M3 S = {{{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}};
  sc = {mod * s.x, mod * s.y, mod * s.z};
  S.m[0][0] = sc.x;
  S.m[1][1] = sc.y;
  S.m[2][2] = sc.z;
  float r = q.x, x = q.y, y = q.z, z = q.w;
  R = {{{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
        {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
        {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}}};
  Mm = mul(S, R);
  M3 Sigma = mul(transpose(Mm), Mm);
  cov3D[0] = Sigma.m[0][0];
  cov3D[1] = Sigma.m[0][1];
  cov3D[2] = Sigma.m[0][2];
  cov3D[3] = Sigma.m[1][1];
  cov3D[4] = Sigma.m[1][2];
  cov3D[5] = Sigma.m[2][2];
}

// ACC: add to the outputs instead of overwriting them (FdlsBwdArgs.accumulate: a rank sums the gradients of its
// views in place, e.g. directly in its slice of the all-reduce bucket; culled Gaussians then write nothing)
template <bool ACC>
__device__ __forceinline__ void put(float* p, float v) {
  *p = ACC ? *p + v : v;
}
template <bool ACC>
__device__ __forceinline__ void put4(float4* p, float4 v) {
  if (ACC) {
    const float4 o = *p;
    v = make_float4(o.x + v.x, o.y + v.y, o.z + v.z, o.w + v.w);
  }
  *p = v;
}

#ifndef FDLS_PB_MINBLOCKS
#define FDLS_PB_MINBLOCKS 3  // 80 registers (~200 B of spills), 3 CTAs per SM: the accumulate variant (3 of 4 views) is latency-bound
#endif
template <bool VEC_SH, bool ACC>
__global__ void __launch_bounds__(256, FDLS_PB_MINBLOCKS)
preprocess_bwd_kernel(int P, int D, int M, int F, const float* __restrict__ means3D, const int* __restrict__ radii,
                      const float* __restrict__ shs, const uint8_t* __restrict__ clamped,
                      const float* __restrict__ scales, const float* __restrict__ rotations,
                      const float* __restrict__ cov3D_precomp, ViewParams vp, const float4* __restrict__ grad_rec,
                      BwdOut o) {
  pdl_grid_sync();
  __shared__ float s_view[16], s_proj[16], s_cam[3];
  if (threadIdx.x < 16) s_view[threadIdx.x] = vp.view[threadIdx.x];
  else if (threadIdx.x < 32) s_proj[threadIdx.x - 16] = vp.proj[threadIdx.x - 16];
  else if (threadIdx.x < 35) s_cam[threadIdx.x - 32] = vp.campos != nullptr ? vp.campos[threadIdx.x - 32] : 0.f;
  __syncthreads();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  // M == 16: this warp's 32 x 192-byte SH block is contiguous.  It is streamed into shared memory with coalesced 16-byte
  // loads (all twelve in flight before the first store) and re-read per thread conflict-free (row stride 49 words); the
  // direct per-thread reads (45 scalar loads at a 192-byte lane stride, backward.cu:20-139 reads sh[idx*max_coeffs + i])
  // touch 32 sectors per instruction for 4 bytes each.
  extern __shared__ float pb_sh[];
  constexpr int PB_SH_STRIDE = 49;
  if (VEC_SH && shs != nullptr && D > 0) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g0 = blockIdx.x * 256 + warp * 32;
    if (g0 < P) {
      const int n_fl = min(32, P - g0) * 48;
      const float* src = shs + (size_t)g0 * 48;
      float* dst = pb_sh + warp * 32 * PB_SH_STRIDE;
      if (n_fl == 1536) {
        const float4* src4 = reinterpret_cast<const float4*>(src);
        float4 q[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) q[i] = __ldg(src4 + lane + 32 * i);
#pragma unroll
        for (int i = 0; i < 12; ++i) {
          const uint32_t j = lane + 32 * i;            // 16-byte group; a row is twelve of them
          const uint32_t g = __umulhi(j, 357913942u);  // j / 12 for j < 2^16
          float* d = dst + g * PB_SH_STRIDE + (j - g * 12) * 4;
          d[0] = q[i].x; d[1] = q[i].y; d[2] = q[i].z; d[3] = q[i].w;
        }
      } else {
        for (int e = lane; e < n_fl; e += 32) dst[(e / 48) * PB_SH_STRIDE + e % 48] = src[e];
      }
    }
    __syncwarp();
  }
  if (idx >= P) return;
  const bool vis = radii[idx] > 0;

  float g[16];
  if (vis) {
    const float4* gr = grad_rec + (size_t)idx * 4;
    const float4 g0 = gr[0], g1 = gr[1], g2 = gr[2], g3 = gr[3];
    g[0] = g0.x; g[1] = g0.y; g[2] = g0.z; g[3] = g0.w; g[4] = g1.x; g[5] = g1.y; g[6] = g1.z; g[7] = g1.w;
    g[8] = g2.x; g[9] = g2.y; g[10] = g2.z; g[11] = g2.w; g[12] = g3.x; g[13] = g3.y; g[14] = g3.z; g[15] = g3.w;
  } else {
#pragma unroll
    for (int k = 0; k < 16; ++k) g[k] = 0.f;
  }
  if (o.sh_rgb && o.sh_rgb_header && idx == 0) {
    o.sh_rgb[-4] = s_cam[0];
    o.sh_rgb[-3] = s_cam[1];
    o.sh_rgb[-2] = s_cam[2];
    o.sh_rgb[-1] = 1.0f;
  }
  if (o.sh_rgb) {  // backward.cu:53-56: clamped channels pass no gradient to the SH coefficients
    const uint8_t cl = vis ? clamped[idx] : (uint8_t)0;
    o.sh_rgb[3 * idx] = (cl & 1) ? 0.f : g[G_RGB];
    o.sh_rgb[3 * idx + 1] = (cl & 2) ? 0.f : g[G_RGB + 1];
    o.sh_rgb[3 * idx + 2] = (cl & 4) ? 0.f : g[G_RGB + 2];
  }
  if (ACC && !vis) return;  // adds nothing
  // pass-through gradients (zeros when culled)
  if (o.means2D) {
    put<ACC>(o.means2D + 3 * idx, g[G_MX]);
    put<ACC>(o.means2D + 3 * idx + 1, g[G_MY]);
    if (!ACC) o.means2D[3 * idx + 2] = 0.f;
  }
  if (o.colors) {
    put<ACC>(o.colors + 3 * idx, g[G_RGB]);
    put<ACC>(o.colors + 3 * idx + 1, g[G_RGB + 1]);
    put<ACC>(o.colors + 3 * idx + 2, g[G_RGB + 2]);
  }
  if (o.lang) {
#pragma unroll
    for (int k = 0; k < FDLS_MAX_LANG_CHANNELS; ++k)
      if (k < F) put<ACC>(o.lang + (size_t)idx * F + k, g[g_lang_slot(k)]);
  }
  if (o.opacity) put<ACC>(o.opacity + idx, g[G_OP]);
  if (o.depths) put<ACC>(o.depths + idx, g[G_DEPTH]);
  if (o.conic) put4<ACC>(reinterpret_cast<float4*>(o.conic) + idx, make_float4(g[G_CA], g[G_CB], 0.f, g[G_CC]));

  // nothing but pass-through gradients wanted (e.g. the fine-lang stage trains the language features only)
  if (!o.means3D && !o.cov3D && !o.sh && !o.scales && !o.rots) return;
  float dL_dcov[6] = {0, 0, 0, 0, 0, 0};
  float3 dL_dmeans = {0, 0, 0};
  float dsc[3] = {0, 0, 0};
  float4 drot = {0, 0, 0, 0};
  const bool want_sh = (shs != nullptr) && (o.sh != nullptr);
  float* sh_out = want_sh ? o.sh + (size_t)idx * M * 3 : nullptr;

  if (!vis) {
    if (want_sh) {
      if (VEC_SH) {
        float4* p = reinterpret_cast<float4*>(sh_out);
        for (int k = 0; k < (M * 3) / 4; ++k) p[k] = make_float4(0, 0, 0, 0);
      } else {
        for (int k = 0; k < M * 3; ++k) sh_out[k] = 0.f;
      }
    }
  } else {
    const float3 mean = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
    float cov3D[6];
    M3 Mm, R;
    float3 sc = {0, 0, 0};
    const bool have_sr = (scales != nullptr);
    if (cov3D_precomp != nullptr) {
#pragma unroll
      for (int i = 0; i < 6; ++i) cov3D[i] = cov3D_precomp[6 * idx + i];
    } else {
      const float3 s = {scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]};
      const float4 q = reinterpret_cast<const float4*>(rotations)[idx];
      cov3d_build(s, vp.scale_modifier, q, cov3D, Mm, R, sc);
    }
    // ---------------- computeCov2DCUDA (backward.cu:144-281)
    {
      const float3 dL_dconic = {g[G_CA], g[G_CB], g[G_CC]};
      float3 t = xform4x3(mean, s_view);
      const float limx = 1.3f * vp.tanfovx;
      const float limy = 1.3f * vp.tanfovy;
      const float txtz = t.x / t.z;
      const float tytz = t.y / t.z;
      t.x = min(limx, max(-limx, txtz)) * t.z;
      t.y = min(limy, max(-limy, tytz)) * t.z;
      const float x_grad_mul = txtz < -limx || txtz > limx ? 0 : 1;
      const float y_grad_mul = tytz < -limy || tytz > limy ? 0 : 1;
      const float h_x = vp.focal_x, h_y = vp.focal_y;
      M3 J = {{{h_x / t.z, 0.0f, -(h_x * t.x) / (t.z * t.z)}, {0.0f, h_y / t.z, -(h_y * t.y) / (t.z * t.z)}, {0, 0, 0}}};
      const float* vm = s_view;
      M3 Wm = {{{vm[0], vm[4], vm[8]}, {vm[1], vm[5], vm[9]}, {vm[2], vm[6], vm[10]}}};
      M3 Vrk = {{{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}}};
      M3 T = mul(Wm, J);
      M3 cov2D = mul(mul(transpose(T), transpose(Vrk)), T);
      const float a = cov2D.m[0][0] += 0.3f;
      const float b = cov2D.m[0][1];
      const float c = cov2D.m[1][1] += 0.3f;
      const float denom = a * c - b * b;
      float dL_da = 0, dL_db = 0, dL_dc = 0;
      const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
      if (denom2inv != 0) {
        dL_da = denom2inv * (-c * c * dL_dconic.x + 2 * b * c * dL_dconic.y + (denom - a * c) * dL_dconic.z);
        dL_dc = denom2inv * (-a * a * dL_dconic.z + 2 * a * b * dL_dconic.y + (denom - a * c) * dL_dconic.x);
        dL_db = denom2inv * 2 * (b * c * dL_dconic.x - (denom + 2 * b * b) * dL_dconic.y + a * b * dL_dconic.z);
        dL_dcov[0] = (T.m[0][0] * T.m[0][0] * dL_da + T.m[0][0] * T.m[1][0] * dL_db + T.m[1][0] * T.m[1][0] * dL_dc);
        dL_dcov[3] = (T.m[0][1] * T.m[0][1] * dL_da + T.m[0][1] * T.m[1][1] * dL_db + T.m[1][1] * T.m[1][1] * dL_dc);
        dL_dcov[5] = (T.m[0][2] * T.m[0][2] * dL_da + T.m[0][2] * T.m[1][2] * dL_db + T.m[1][2] * T.m[1][2] * dL_dc);
        dL_dcov[1] = 2 * T.m[0][0] * T.m[0][1] * dL_da + (T.m[0][0] * T.m[1][1] + T.m[0][1] * T.m[1][0]) * dL_db +
                     2 * T.m[1][0] * T.m[1][1] * dL_dc;
        dL_dcov[2] = 2 * T.m[0][0] * T.m[0][2] * dL_da + (T.m[0][0] * T.m[1][2] + T.m[0][2] * T.m[1][0]) * dL_db +
                     2 * T.m[1][0] * T.m[1][2] * dL_dc;
        dL_dcov[4] = 2 * T.m[0][2] * T.m[0][1] * dL_da + (T.m[0][1] * T.m[1][2] + T.m[0][2] * T.m[1][1]) * dL_db +
                     2 * T.m[1][1] * T.m[1][2] * dL_dc;
      }
      const float dL_dT00 = 2 * (T.m[0][0] * Vrk.m[0][0] + T.m[0][1] * Vrk.m[0][1] + T.m[0][2] * Vrk.m[0][2]) * dL_da +
                            (T.m[1][0] * Vrk.m[0][0] + T.m[1][1] * Vrk.m[0][1] + T.m[1][2] * Vrk.m[0][2]) * dL_db;
      const float dL_dT01 = 2 * (T.m[0][0] * Vrk.m[1][0] + T.m[0][1] * Vrk.m[1][1] + T.m[0][2] * Vrk.m[1][2]) * dL_da +
                            (T.m[1][0] * Vrk.m[1][0] + T.m[1][1] * Vrk.m[1][1] + T.m[1][2] * Vrk.m[1][2]) * dL_db;
      const float dL_dT02 = 2 * (T.m[0][0] * Vrk.m[2][0] + T.m[0][1] * Vrk.m[2][1] + T.m[0][2] * Vrk.m[2][2]) * dL_da +
                            (T.m[1][0] * Vrk.m[2][0] + T.m[1][1] * Vrk.m[2][1] + T.m[1][2] * Vrk.m[2][2]) * dL_db;
      const float dL_dT10 = 2 * (T.m[1][0] * Vrk.m[0][0] + T.m[1][1] * Vrk.m[0][1] + T.m[1][2] * Vrk.m[0][2]) * dL_dc +
                            (T.m[0][0] * Vrk.m[0][0] + T.m[0][1] * Vrk.m[0][1] + T.m[0][2] * Vrk.m[0][2]) * dL_db;
      const float dL_dT11 = 2 * (T.m[1][0] * Vrk.m[1][0] + T.m[1][1] * Vrk.m[1][1] + T.m[1][2] * Vrk.m[1][2]) * dL_dc +
                            (T.m[0][0] * Vrk.m[1][0] + T.m[0][1] * Vrk.m[1][1] + T.m[0][2] * Vrk.m[1][2]) * dL_db;
      const float dL_dT12 = 2 * (T.m[1][0] * Vrk.m[2][0] + T.m[1][1] * Vrk.m[2][1] + T.m[1][2] * Vrk.m[2][2]) * dL_dc +
                            (T.m[0][0] * Vrk.m[2][0] + T.m[0][1] * Vrk.m[2][1] + T.m[0][2] * Vrk.m[2][2]) * dL_db;
      const float dL_dJ00 = Wm.m[0][0] * dL_dT00 + Wm.m[0][1] * dL_dT01 + Wm.m[0][2] * dL_dT02;
      const float dL_dJ02 = Wm.m[2][0] * dL_dT00 + Wm.m[2][1] * dL_dT01 + Wm.m[2][2] * dL_dT02;
      const float dL_dJ11 = Wm.m[1][0] * dL_dT10 + Wm.m[1][1] * dL_dT11 + Wm.m[1][2] * dL_dT12;
      const float dL_dJ12 = Wm.m[2][0] * dL_dT10 + Wm.m[2][1] * dL_dT11 + Wm.m[2][2] * dL_dT12;
      const float tz = 1.f / t.z;
      const float tz2 = tz * tz;
      const float tz3 = tz2 * tz;
      const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
      const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
      const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 +
                           (2 * h_y * t.y) * tz3 * dL_dJ12;
      dL_dmeans = xform_vec4x3_T({dL_dtx, dL_dty, dL_dtz}, s_view);
    }
    // ---------------- preprocessCUDA backward (backward.cu:353-431)
    {
      const float3 m = mean;
      const float* proj = s_proj;
      const float* view = s_view;
      const float4 m_hom = xform4x4(m, proj);
      const float m_w = 1.0f / (m_hom.w + 0.0000001f);
      const float dmx = g[G_MX], dmy = g[G_MY];
      const float mul1 = (proj[0] * m.x + proj[4] * m.y + proj[8] * m.z + proj[12]) * m_w * m_w;
      const float mul2 = (proj[1] * m.x + proj[5] * m.y + proj[9] * m.z + proj[13]) * m_w * m_w;
      V3 dL_dmean;
      dL_dmean.x = (proj[0] * m_w - proj[3] * mul1) * dmx + (proj[1] * m_w - proj[3] * mul2) * dmy;
      dL_dmean.y = (proj[4] * m_w - proj[7] * mul1) * dmx + (proj[5] * m_w - proj[7] * mul2) * dmy;
      dL_dmean.z = (proj[8] * m_w - proj[11] * mul1) * dmx + (proj[9] * m_w - proj[11] * mul2) * dmy;
      dL_dmeans.x += dL_dmean.x;
      dL_dmeans.y += dL_dmean.y;
      dL_dmeans.z += dL_dmean.z;
      const float dd = g[G_DEPTH];
      const float mul3 = view[2] * m.x + view[6] * m.y + view[10] * m.z + view[14];
      V3 dL_dmean2;
      dL_dmean2.x = (view[2] - view[3] * mul3) * dd;
      dL_dmean2.y = (view[6] - view[7] * mul3) * dd;
      dL_dmean2.z = (view[10] - view[11] * mul3) * dd;
      dL_dmeans.x += dL_dmean2.x;
      dL_dmeans.y += dL_dmean2.y;
      dL_dmeans.z += dL_dmean2.z;
    }
    // ---------------- SH backward (backward.cu:20-139)
    if (shs != nullptr) {
      const float* shp = VEC_SH ? pb_sh + threadIdx.x * PB_SH_STRIDE : shs + (size_t)idx * M * 3;
      auto SH = [&](int i) { return V3{shp[3 * i], shp[3 * i + 1], shp[3 * i + 2]}; };
      V3 pos = {mean.x, mean.y, mean.z};
      V3 cam = {s_cam[0], s_cam[1], s_cam[2]};
      V3 dir_orig = pos - cam;
      V3 dir = dir_orig / sqrt(dot3(dir_orig, dir_orig));
      const uint8_t cl = clamped[idx];
      V3 dL_dRGB = {g[G_RGB], g[G_RGB + 1], g[G_RGB + 2]};
      dL_dRGB.x *= (cl & 1) ? 0 : 1;
      dL_dRGB.y *= (cl & 2) ? 0 : 1;
      dL_dRGB.z *= (cl & 4) ? 0 : 1;
      V3 dRGBdx = {0, 0, 0}, dRGBdy = {0, 0, 0}, dRGBdz = {0, 0, 0};
      const float x = dir.x, y = dir.y, z = dir.z;
      V3 dsh[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) dsh[k] = {0, 0, 0};
      dsh[0] = SH_C0 * dL_dRGB;
      if (D > 0) {
        dsh[1] = (-SH_C1 * y) * dL_dRGB;
        dsh[2] = (SH_C1 * z) * dL_dRGB;
        dsh[3] = (-SH_C1 * x) * dL_dRGB;
        dRGBdx = -SH_C1 * SH(3);
        dRGBdy = -SH_C1 * SH(1);
        dRGBdz = SH_C1 * SH(2);
        if (D > 1) {
          const float xx = x * x, yy = y * y, zz = z * z;
          const float xy = x * y, yz = y * z, xz = x * z;
          dsh[4] = (SH_C2[0] * xy) * dL_dRGB;
          dsh[5] = (SH_C2[1] * yz) * dL_dRGB;
          dsh[6] = (SH_C2[2] * (2.f * zz - xx - yy)) * dL_dRGB;
          dsh[7] = (SH_C2[3] * xz) * dL_dRGB;
          dsh[8] = (SH_C2[4] * (xx - yy)) * dL_dRGB;
          dRGBdx += SH_C2[0] * y * SH(4) + SH_C2[2] * 2.f * -x * SH(6) + SH_C2[3] * z * SH(7) + SH_C2[4] * 2.f * x * SH(8);
          dRGBdy += SH_C2[0] * x * SH(4) + SH_C2[1] * z * SH(5) + SH_C2[2] * 2.f * -y * SH(6) + SH_C2[4] * 2.f * -y * SH(8);
          dRGBdz += SH_C2[1] * y * SH(5) + SH_C2[2] * 2.f * 2.f * z * SH(6) + SH_C2[3] * x * SH(7);
          if (D > 2) {
            dsh[9] = (SH_C3[0] * y * (3.f * xx - yy)) * dL_dRGB;
            dsh[10] = (SH_C3[1] * xy * z) * dL_dRGB;
            dsh[11] = (SH_C3[2] * y * (4.f * zz - xx - yy)) * dL_dRGB;
            dsh[12] = (SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy)) * dL_dRGB;
            dsh[13] = (SH_C3[4] * x * (4.f * zz - xx - yy)) * dL_dRGB;
            dsh[14] = (SH_C3[5] * z * (xx - yy)) * dL_dRGB;
            dsh[15] = (SH_C3[6] * x * (xx - 3.f * yy)) * dL_dRGB;
            dRGBdx += (SH_C3[0] * SH(9) * 3.f * 2.f * xy + SH_C3[1] * SH(10) * yz + SH_C3[2] * SH(11) * -2.f * xy +
                       SH_C3[3] * SH(12) * -3.f * 2.f * xz + SH_C3[4] * SH(13) * (-3.f * xx + 4.f * zz - yy) +
                       SH_C3[5] * SH(14) * 2.f * xz + SH_C3[6] * SH(15) * 3.f * (xx - yy));
            dRGBdy += (SH_C3[0] * SH(9) * 3.f * (xx - yy) + SH_C3[1] * SH(10) * xz +
                       SH_C3[2] * SH(11) * (-3.f * yy + 4.f * zz - xx) + SH_C3[3] * SH(12) * -3.f * 2.f * yz +
                       SH_C3[4] * SH(13) * -2.f * xy + SH_C3[5] * SH(14) * -2.f * yz +
                       SH_C3[6] * SH(15) * -3.f * 2.f * xy);
            dRGBdz += (SH_C3[1] * SH(10) * xy + SH_C3[2] * SH(11) * 4.f * 2.f * yz +
                       SH_C3[3] * SH(12) * 3.f * (2.f * zz - xx - yy) + SH_C3[4] * SH(13) * 4.f * 2.f * xz +
                       SH_C3[5] * SH(14) * (xx - yy));
          }
        }
      }
      if (want_sh) {
        if (VEC_SH) {  // M == 16: 12 x 16-byte stores
          float4* p = reinterpret_cast<float4*>(sh_out);
          const float* f = reinterpret_cast<const float*>(dsh);
#pragma unroll
          for (int k = 0; k < 12; ++k) put4<ACC>(p + k, make_float4(f[4 * k], f[4 * k + 1], f[4 * k + 2], f[4 * k + 3]));
        } else {
#pragma unroll
          for (int k = 0; k < 16; ++k)
            if (k < M) {
              put<ACC>(sh_out + 3 * k, dsh[k].x);
              put<ACC>(sh_out + 3 * k + 1, dsh[k].y);
              put<ACC>(sh_out + 3 * k + 2, dsh[k].z);
            }
        }
      }
      const float3 dL_ddir = {dot3(dRGBdx, dL_dRGB), dot3(dRGBdy, dL_dRGB), dot3(dRGBdz, dL_dRGB)};
      const float3 dm = dnormvdv(float3{dir_orig.x, dir_orig.y, dir_orig.z}, dL_ddir);
      dL_dmeans.x += dm.x;
      dL_dmeans.y += dm.y;
      dL_dmeans.z += dm.z;
    }
    // ---------------- cov3D backward (backward.cu:285-348)
    if (have_sr && cov3D_precomp == nullptr) {
      const float4 q = reinterpret_cast<const float4*>(rotations)[idx];
      const float r = q.x, x = q.y, y = q.z, z = q.w;
      M3 dL_dSigma = {{{dL_dcov[0], 0.5f * dL_dcov[1], 0.5f * dL_dcov[2]},
                       {0.5f * dL_dcov[1], dL_dcov[3], 0.5f * dL_dcov[4]},
                       {0.5f * dL_dcov[2], 0.5f * dL_dcov[4], dL_dcov[5]}}};
      M3 M2;
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int k = 0; k < 3; ++k) M2.m[c][k] = 2.0f * Mm.m[c][k];
      M3 dL_dM = mul(M2, dL_dSigma);
      M3 Rt = transpose(R);
      M3 dL_dMt = transpose(dL_dM);
      auto col = [](const M3& A, int c) { return V3{A.m[c][0], A.m[c][1], A.m[c][2]}; };
      dsc[0] = dot3(col(Rt, 0), col(dL_dMt, 0));
      dsc[1] = dot3(col(Rt, 1), col(dL_dMt, 1));
      dsc[2] = dot3(col(Rt, 2), col(dL_dMt, 2));
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        dL_dMt.m[0][k] *= sc.x;
        dL_dMt.m[1][k] *= sc.y;
        dL_dMt.m[2][k] *= sc.z;
      }
      const auto& t = dL_dMt.m;
      drot.x = 2 * z * (t[0][1] - t[1][0]) + 2 * y * (t[2][0] - t[0][2]) + 2 * x * (t[1][2] - t[2][1]);
      drot.y = 2 * y * (t[1][0] + t[0][1]) + 2 * z * (t[2][0] + t[0][2]) + 2 * r * (t[1][2] - t[2][1]) -
               4 * x * (t[2][2] + t[1][1]);
      drot.z = 2 * x * (t[1][0] + t[0][1]) + 2 * r * (t[2][0] - t[0][2]) + 2 * z * (t[1][2] + t[2][1]) -
               4 * y * (t[2][2] + t[0][0]);
      drot.w = 2 * r * (t[0][1] - t[1][0]) + 2 * x * (t[2][0] + t[0][2]) + 2 * y * (t[1][2] + t[2][1]) -
               4 * z * (t[1][1] + t[0][0]);
    }
  }
  if (o.means3D) {
    put<ACC>(o.means3D + 3 * idx, dL_dmeans.x);
    put<ACC>(o.means3D + 3 * idx + 1, dL_dmeans.y);
    put<ACC>(o.means3D + 3 * idx + 2, dL_dmeans.z);
  }
  if (o.cov3D) {
#pragma unroll
    for (int i = 0; i < 6; ++i) put<ACC>(o.cov3D + 6 * idx + i, dL_dcov[i]);
  }
  if (o.scales) {
    put<ACC>(o.scales + 3 * idx, dsc[0]);
    put<ACC>(o.scales + 3 * idx + 1, dsc[1]);
    put<ACC>(o.scales + 3 * idx + 2, dsc[2]);
  }
  if (o.rots) put4<ACC>(reinterpret_cast<float4*>(o.rots) + idx, drot);
}

// dL_dsh = sum over views of basis(dir_v) (x) dL_dRGB_v: the SH gradient of one view is an outer product of the 16
// view-dependent basis values (backward.cu:41-126) and the masked colour gradient, so a batch of B views needs
// 3 floats per view and Gaussian, not 48 per Gaussian and rank, to cross the interconnect (dist.GradBucket lowrank_sh).
// slots: B records of `stride` floats: [0..2] camera position, [3] 1.0 if the slot holds a view, [4..] (P,3) dL_dRGB.
constexpr int SGE_THREADS = 128;  // 128 x 49 words of staging = 25 KB of static shared memory
__global__ void __launch_bounds__(SGE_THREADS)
sh_grad_expand_kernel(int P, int D, int M, int B, const float* __restrict__ means3D, const float* __restrict__ slots,
                      size_t stride, float* __restrict__ dL_dsh, int accumulate) {
  pdl_grid_sync();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const V3 pos = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
  V3 dsh[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) dsh[k] = {0, 0, 0};
  for (int v = 0; v < B; ++v) {
    const float* sl = slots + (size_t)v * stride;
    if (sl[3] == 0.f) continue;
    const V3 g = {sl[4 + 3 * idx], sl[4 + 3 * idx + 1], sl[4 + 3 * idx + 2]};
    if (g.x == 0.f && g.y == 0.f && g.z == 0.f) continue;  // culled in this view (its direction may be degenerate)
    const V3 cam = {sl[0], sl[1], sl[2]};
    const V3 dir_orig = pos - cam;
    const V3 dir = dir_orig / sqrt(dot3(dir_orig, dir_orig));
    const float x = dir.x, y = dir.y, z = dir.z;
    dsh[0] += SH_C0 * g;
    if (D > 0) {
      dsh[1] += (-SH_C1 * y) * g;
      dsh[2] += (SH_C1 * z) * g;
      dsh[3] += (-SH_C1 * x) * g;
      if (D > 1) {
        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        dsh[4] += (SH_C2[0] * xy) * g;
        dsh[5] += (SH_C2[1] * yz) * g;
        dsh[6] += (SH_C2[2] * (2.f * zz - xx - yy)) * g;
        dsh[7] += (SH_C2[3] * xz) * g;
        dsh[8] += (SH_C2[4] * (xx - yy)) * g;
        if (D > 2) {
          dsh[9] += (SH_C3[0] * y * (3.f * xx - yy)) * g;
          dsh[10] += (SH_C3[1] * xy * z) * g;
          dsh[11] += (SH_C3[2] * y * (4.f * zz - xx - yy)) * g;
          dsh[12] += (SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy)) * g;
          dsh[13] += (SH_C3[4] * x * (4.f * zz - xx - yy)) * g;
          dsh[14] += (SH_C3[5] * z * (xx - yy)) * g;
          dsh[15] += (SH_C3[6] * x * (xx - 3.f * yy)) * g;
        }
      }
    }
  }
  float* out = dL_dsh + (size_t)idx * M * 3;
  if (M == 16) {
    // A warp's 32 rows of 192 bytes are contiguous: they go through shared memory (row stride 49 words, conflict-free
    // both ways) and leave as coalesced 16-byte stores, instead of twelve 16-byte stores per thread at a 192-byte lane
    // stride (32 sectors per instruction).  The last, partial warp of the array stores directly.
    __shared__ float s_out[SGE_THREADS * 49];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g0 = blockIdx.x * SGE_THREADS + warp * 32;
    if (g0 + 32 <= P) {
      float* row = s_out + threadIdx.x * 49;
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        row[3 * k] = dsh[k].x;
        row[3 * k + 1] = dsh[k].y;
        row[3 * k + 2] = dsh[k].z;
      }
      __syncwarp();
      float4* dst = reinterpret_cast<float4*>(dL_dsh + (size_t)g0 * 48);
      const float* src = s_out + warp * 32 * 49;
#pragma unroll
      for (int i = 0; i < 12; ++i) {
        const uint32_t j = lane + 32 * i;            // 16-byte group of the warp's block; a row is twelve of them
        const uint32_t g = __umulhi(j, 357913942u);  // j / 12
        const float* q = src + g * 49 + (j - g * 12) * 4;
        float4 v = make_float4(q[0], q[1], q[2], q[3]);
        if (accumulate) {
          const float4 c = dst[j];
          v = make_float4(c.x + v.x, c.y + v.y, c.z + v.z, c.w + v.w);
        }
        dst[j] = v;
      }
      return;
    }
    float4* o4 = reinterpret_cast<float4*>(out);
    const float* f = reinterpret_cast<const float*>(dsh);
#pragma unroll
    for (int k = 0; k < 12; ++k) {
      float4 v = make_float4(f[4 * k], f[4 * k + 1], f[4 * k + 2], f[4 * k + 3]);
      if (accumulate) {
        const float4 c = o4[k];
        v = make_float4(c.x + v.x, c.y + v.y, c.z + v.z, c.w + v.w);
      }
      o4[k] = v;
    }
    return;
  }
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    if (k >= M) break;
    if (accumulate) {
      out[3 * k] += dsh[k].x; out[3 * k + 1] += dsh[k].y; out[3 * k + 2] += dsh[k].z;
    } else {
      out[3 * k] = dsh[k].x; out[3 * k + 1] = dsh[k].y; out[3 * k + 2] = dsh[k].z;
    }
  }
  for (int k = 16; k < M; ++k)
    if (!accumulate) out[3 * k] = out[3 * k + 1] = out[3 * k + 2] = 0.f;
}

cudaError_t launch_sh_grad_expand(int P, int D, int M, int B, const float* means3D, const float* slots, size_t stride,
                                  float* dL_dsh, int accumulate, cudaStream_t st) {
  if (P == 0 || B == 0) return cudaSuccess;
  count_launch();
  launch_pdl(sh_grad_expand_kernel, dim3(cdiv(P, SGE_THREADS)), dim3(SGE_THREADS), 0, st, P, D, M, B, means3D, slots, stride, dL_dsh,
             accumulate);
  return cudaGetLastError();
}

cudaError_t launch_preprocess_bwd(const FdlsBwdArgs* a, const ViewParams& vp, const uint8_t* clamped,
                                  const float4* grad_rec, cudaStream_t st) {
  if (a->P == 0) return cudaSuccess;
  BwdOut o = {a->dL_dmeans3D, a->dL_dmeans2D, a->dL_dsh, a->dL_dcolors, a->dL_dlanguage, a->dL_dopacity,
              a->dL_dscales, a->dL_drotations, a->dL_dcov3D, a->dL_dconic, a->dL_ddepths, a->dL_dsh_rgb,
              a->sh_rgb_header};
  const int grid = cdiv(a->P, 256);
  const size_t pb_smem = (size_t)256 * 49 * sizeof(float);  // staged SH rows (M == 16)
  count_launch();
#define PB_LAUNCH(VEC, ACCUM)                                                                                      \
  launch_pdl(preprocess_bwd_kernel<VEC, ACCUM>, dim3(grid), dim3(256), VEC ? pb_smem : 0, st, a->P, a->D, a->M, a->F, \
             a->means3D, a->radii, a->shs, clamped, a->scales, a->rotations, a->cov3D_precomp, vp, grad_rec, o)
  if (a->M == 16) {
    static SmemOptIn opt_acc, opt_set;  // per device (common.cuh)
    if (a->accumulate) {
      if (cudaError_t e = opt_acc.ensure(preprocess_bwd_kernel<true, true>, pb_smem)) return e;
      PB_LAUNCH(true, true);
    } else {
      if (cudaError_t e = opt_set.ensure(preprocess_bwd_kernel<true, false>, pb_smem)) return e;
      PB_LAUNCH(true, false);
    }
  } else {
    if (a->accumulate) PB_LAUNCH(false, true); else PB_LAUNCH(false, false);
  }
#undef PB_LAUNCH
  return cudaGetLastError();
}

}  // namespace fdls
