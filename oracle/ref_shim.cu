// TEST INFRASTRUCTURE ONLY -- C-ABI shim around the UNMODIFIED reference CUDA rasterizer.
//
// oracle/build_ref.py compiles this file together with the reference's own three translation units
//   /root/reference/submodules/4d-langsplat-rasterization/cuda_rasterizer/{forward,backward,rasterizer_impl}.cu
// (from where they lie; nothing is copied into this repository) into oracle/_ref/libref_f{3,6}.so for sm_100a.
// The shim replaces the torch glue of the reference (rasterize_points.cu:35-235), which only allocates tensors
// and forwards raw pointers to CudaRasterizer::Rasterizer::{forward,backward,markVisible} (rasterizer.h:20-95).
// Only tests/, __graft_entry__.smoke() and bench.py's reference arm may load the resulting library.
#include <cstdint>
#include <cstdio>
#include <functional>
#include <cuda_runtime.h>
#include "rasterizer_impl.h"  // reference header, found through -I at build time

namespace {

struct Buf {
  char* p = nullptr;
  size_t cap = 0;
  char* resize(size_t n) {  // the reference's resizeFunctional (rasterize_points.cu:27-33) without torch
    if (n > cap) {
      if (p) cudaFree(p);
      size_t want = n + n / 4 + 256;
      if (cudaMalloc(&p, want) != cudaSuccess) { p = nullptr; cap = 0; return nullptr; }
      cap = want;
    }
    return p;
  }
  ~Buf() { if (p) cudaFree(p); }
};

struct Handle {
  Buf geom, binning, img;
};

}  // namespace

extern "C" {

struct RefState {  // device pointers into the three reference buffers (rasterizer_impl.cu:155-194)
  float* depths; bool* clamped; int* internal_radii; float2* means2D; float* cov3D; float4* conic_opacity;
  float* rgb; uint32_t* point_offsets; uint32_t* tiles_touched;
  uint2* ranges; uint32_t* n_contrib; float* accum_alpha;
  uint64_t* point_list_keys_unsorted; uint64_t* point_list_keys; uint32_t* point_list_unsorted; uint32_t* point_list;
};

__attribute__((visibility("default"))) void* ref_create() { return new Handle(); }
__attribute__((visibility("default"))) void ref_destroy(void* h) { delete static_cast<Handle*>(h); }
__attribute__((visibility("default"))) int ref_lang_channels() { return NUM_CHANNELS_language_feature; }

__attribute__((visibility("default"))) int ref_forward(
    void* hv, int P, int D, int M, const float* bg, int W, int H, const float* means3D, const float* shs,
    const float* colors_precomp, const float* lang, const float* opacities, const float* scales, float scale_modifier,
    const float* rotations, const float* cov3D_precomp, const float* view, const float* proj, const float* campos,
    float tanfovx, float tanfovy, int prefiltered, float* out_color, float* out_lang, float* out_depth, int* radii,
    int debug, int include_feature) {
  Handle* h = static_cast<Handle*>(hv);
  std::function<char*(size_t)> g = [h](size_t n) { return h->geom.resize(n); };
  std::function<char*(size_t)> b = [h](size_t n) { return h->binning.resize(n); };
  std::function<char*(size_t)> i = [h](size_t n) { return h->img.resize(n); };
  try {
    return CudaRasterizer::Rasterizer::forward(g, b, i, P, D, M, bg, W, H, means3D, shs, colors_precomp, lang,
                                               opacities, scales, scale_modifier, rotations, cov3D_precomp, view, proj,
                                               campos, tanfovx, tanfovy, prefiltered != 0, out_color, out_lang,
                                               out_depth, radii, debug != 0, include_feature != 0);
  } catch (const std::exception& e) {
    fprintf(stderr, "ref_forward: %s\n", e.what());
    return -1;
  }
}

__attribute__((visibility("default"))) int ref_backward(
    void* hv, int P, int D, int M, int R, const float* bg, int W, int H, const float* means3D, const float* shs,
    const float* colors_precomp, const float* lang, const float* scales, float scale_modifier, const float* rotations,
    const float* cov3D_precomp, const float* view, const float* proj, const float* campos, float tanfovx,
    float tanfovy, const int* radii, const float* dL_dpix, const float* dL_dpix_F, const float* dL_dpix_depth,
    float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_dlang, float* dL_ddepth,
    float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, int debug,
    int include_feature) {
  Handle* h = static_cast<Handle*>(hv);
  try {
    CudaRasterizer::Rasterizer::backward(P, D, M, R, bg, W, H, means3D, shs, colors_precomp, lang, scales,
                                         scale_modifier, rotations, cov3D_precomp, view, proj, campos, tanfovx, tanfovy,
                                         radii, h->geom.p, h->binning.p, h->img.p, dL_dpix, dL_dpix_F, dL_dpix_depth,
                                         dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dlang, dL_ddepth, dL_dmean3D,
                                         dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug != 0, include_feature != 0);
  } catch (const std::exception& e) {
    fprintf(stderr, "ref_backward: %s\n", e.what());
    return -1;
  }
  return 0;
}

__attribute__((visibility("default"))) void ref_mark_visible(int P, float* means3D, float* view, float* proj,
                                                             bool* present) {
  CudaRasterizer::Rasterizer::markVisible(P, means3D, view, proj, present);
}

// Plain device-to-device copy so the tests can pull intermediates out of the reference buffers.
__attribute__((visibility("default"))) int ref_copy_d2d(void* dst, const void* src, size_t bytes) {
  return (int)cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToDevice);
}

// Decode the three byte buffers of the last forward with the reference's own fromChunk layouts.
__attribute__((visibility("default"))) void ref_state(void* hv, int P, int R, int HW, RefState* o) {
  Handle* h = static_cast<Handle*>(hv);
  char* c = h->geom.p;
  CudaRasterizer::GeometryState g = CudaRasterizer::GeometryState::fromChunk(c, P);
  o->depths = g.depths; o->clamped = g.clamped; o->internal_radii = g.internal_radii; o->means2D = g.means2D;
  o->cov3D = g.cov3D; o->conic_opacity = g.conic_opacity; o->rgb = g.rgb; o->point_offsets = g.point_offsets;
  o->tiles_touched = g.tiles_touched;
  c = h->img.p;
  CudaRasterizer::ImageState im = CudaRasterizer::ImageState::fromChunk(c, HW);
  o->ranges = im.ranges; o->n_contrib = im.n_contrib; o->accum_alpha = im.accum_alpha;
  c = h->binning.p;
  if (c && R > 0) {
    CudaRasterizer::BinningState b = CudaRasterizer::BinningState::fromChunk(c, R);
    o->point_list_keys_unsorted = b.point_list_keys_unsorted; o->point_list_keys = b.point_list_keys;
    o->point_list_unsorted = b.point_list_unsorted; o->point_list = b.point_list;
  } else {
    o->point_list_keys_unsorted = nullptr; o->point_list_keys = nullptr; o->point_list_unsorted = nullptr;
    o->point_list = nullptr;
  }
}

}  // extern "C"
