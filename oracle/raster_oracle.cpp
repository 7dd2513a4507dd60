// TEST INFRASTRUCTURE ONLY -- CPU restatement ("oracle") of the 4D LangSplat rasterizer hot path.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
// this library.  The product path (4dlangsplat_b200/) never links, imports or calls it.
//
// Parity status: pinned against the reference CUDA rasterizer rebuilt for sm_100a (oracle/_ref, built by
// oracle/build_ref.py) through the fixtures under tests/golden/ (generated on a B200 by
// tests/golden/make_golden.py).  The reference repository itself holds no tests or golden vectors.
//
// Every function cites the reference file:line it restates (paths relative to
// /root/reference/submodules/4d-langsplat-rasterization/).  The code is templated on the real type so
// that a float64 twin exists for finite-difference gradient checks.
//
// Matrix convention: `Mat3` is column-major (m[c][r]) with the usual column-vector product, which is the
// convention of the header-only math library the reference device code is written against.

#include <cmath>
#include <cstdint>
#include <cstring>
#include <algorithm>
#include <vector>
#include <numeric>

#define ORC_API extern "C" __attribute__((visibility("default")))

namespace {

constexpr int TILE = 16;  // cuda_rasterizer/config.h:18-19 (BLOCK_X = BLOCK_Y = 16)

template <typename R> struct V3 { R x, y, z; };

template <typename R> struct Mat3 {
  R m[3][3];  // m[col][row]
};

template <typename R> Mat3<R> mul(const Mat3<R>& a, const Mat3<R>& b) {
  Mat3<R> o;
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) o.m[c][r] = a.m[0][r] * b.m[c][0] + a.m[1][r] * b.m[c][1] + a.m[2][r] * b.m[c][2];
  return o;
}
template <typename R> Mat3<R> transpose(const Mat3<R>& a) {
  Mat3<R> o;
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) o.m[c][r] = a.m[r][c];
  return o;
}

// auxiliary.h:58-77 -- matrices are read in the "transposed" (row-vector) layout: x' = m[0]x+m[4]y+m[8]z+m[12].
template <typename R> V3<R> xform4x3(const V3<R>& p, const R* m) {
  return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
          m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
}
template <typename R> R xform_w(const V3<R>& p, const R* m) { return m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]; }

// auxiliary.h:22-39
const double C0 = 0.28209479177387814, C1 = 0.4886025119029199;
const double C2[5] = {1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792,
                      0.5462742152960396};
const double C3[7] = {-0.5900435899266435, 2.890611442640554,  -0.4570457994644658, 0.3731763325901154,
                      -0.4570457994644658, 1.445305721320277,  -0.5900435899266435};

// auxiliary.h:41-44: evaluated in double because of the double literals, then rounded to the real type.
template <typename R> R ndc2pix(R v, int S) { return (R)((((double)v + 1.0) * S - 1.0) * 0.5); }

// auxiliary.h:46-56: max_radius is an int; the grid clamps are unsigned mins of non-negative ints.
template <typename R>
void get_rect(R px, R py, int rad, int gx, int gy, int& x0, int& y0, int& x1, int& y1) {
  x0 = std::min(gx, std::max(0, (int)((px - rad) / TILE)));
  y0 = std::min(gy, std::max(0, (int)((py - rad) / TILE)));
  x1 = std::min(gx, std::max(0, (int)((px + rad + TILE - 1) / TILE)));
  y1 = std::min(gy, std::max(0, (int)((py + rad + TILE - 1) / TILE)));
}

// forward.cu:118-152 (quaternion deliberately NOT normalised, :127)
template <typename R> void rot_scale_M(const R* s, R mod, const R* q, Mat3<R>& M, Mat3<R>& Rm, R sc[3]) {
  R r = q[0], x = q[1], y = q[2], z = q[3];
  R cols[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - r * z),     2 * (x * z + r * y),
               2 * (x * y + r * z),     1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
               2 * (x * z - r * y),     2 * (y * z + r * x),     1 - 2 * (x * x + y * y)};
  for (int c = 0; c < 3; ++c)
    for (int rr = 0; rr < 3; ++rr) Rm.m[c][rr] = cols[3 * c + rr];
  Mat3<R> S = {};
  sc[0] = mod * s[0]; sc[1] = mod * s[1]; sc[2] = mod * s[2];
  S.m[0][0] = sc[0]; S.m[1][1] = sc[1]; S.m[2][2] = sc[2];
  M = mul(S, Rm);
}
template <typename R> void cov3d_from_scale_rot(const R* s, R mod, const R* q, R* cov) {
  Mat3<R> M, Rm; R sc[3];
  rot_scale_M(s, mod, q, M, Rm, sc);
  Mat3<R> Sg = mul(transpose(M), M);
  cov[0] = Sg.m[0][0]; cov[1] = Sg.m[0][1]; cov[2] = Sg.m[0][2];
  cov[3] = Sg.m[1][1]; cov[4] = Sg.m[1][2]; cov[5] = Sg.m[2][2];
}

// Shared by forward.cu:74-113 and backward.cu:160-202: T = W*J and the clamped view-space mean.
template <typename R>
void ewa_T(const V3<R>& mean, R fx, R fy, R tanx, R tany, const R* view, Mat3<R>& T, Mat3<R>& W, V3<R>& t,
           R& xmul, R& ymul) {
  t = xform4x3(mean, view);
  const R limx = (R)1.3f * tanx, limy = (R)1.3f * tany;
  const R txtz = t.x / t.z, tytz = t.y / t.z;
  t.x = std::min(limx, std::max(-limx, txtz)) * t.z;
  t.y = std::min(limy, std::max(-limy, tytz)) * t.z;
  xmul = (txtz < -limx || txtz > limx) ? 0 : 1;
  ymul = (tytz < -limy || tytz > limy) ? 0 : 1;
  Mat3<R> J = {};
  J.m[0][0] = fx / t.z; J.m[0][2] = -(fx * t.x) / (t.z * t.z);
  J.m[1][1] = fy / t.z; J.m[1][2] = -(fy * t.y) / (t.z * t.z);
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) W.m[c][r] = view[4 * r + c];
  T = mul(W, J);
}
template <typename R> Mat3<R> sym3(const R* c) {
  Mat3<R> V;
  V.m[0][0] = c[0]; V.m[0][1] = c[1]; V.m[0][2] = c[2];
  V.m[1][0] = c[1]; V.m[1][1] = c[3]; V.m[1][2] = c[4];
  V.m[2][0] = c[2]; V.m[2][1] = c[4]; V.m[2][2] = c[5];
  return V;
}

// forward.cu:20-71
template <typename R>
void sh_to_rgb(int deg, int M, const V3<R>& pos, const R* campos, const R* sh, R rgb[3], uint8_t clamped[3]) {
  R dx = pos.x - campos[0], dy = pos.y - campos[1], dz = pos.z - campos[2];
  R len = std::sqrt(dx * dx + dy * dy + dz * dz);
  R x = dx / len, y = dy / len, z = dz / len;
  for (int c = 0; c < 3; ++c) {
    auto S = [&](int k) { return sh[3 * k + c]; };
    R res = (R)C0 * S(0);
    if (deg > 0) {
      res = res - (R)C1 * y * S(1) + (R)C1 * z * S(2) - (R)C1 * x * S(3);
      if (deg > 1) {
        R xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        res = res + (R)C2[0] * xy * S(4) + (R)C2[1] * yz * S(5) + (R)C2[2] * (2 * zz - xx - yy) * S(6) +
              (R)C2[3] * xz * S(7) + (R)C2[4] * (xx - yy) * S(8);
        if (deg > 2) {
          res = res + (R)C3[0] * y * (3 * xx - yy) * S(9) + (R)C3[1] * xy * z * S(10) +
                (R)C3[2] * y * (4 * zz - xx - yy) * S(11) + (R)C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * S(12) +
                (R)C3[4] * x * (4 * zz - xx - yy) * S(13) + (R)C3[5] * z * (xx - yy) * S(14) +
                (R)C3[6] * x * (xx - 3 * yy) * S(15);
        }
      }
    }
    res += (R)0.5;
    clamped[c] = res < 0;
    rgb[c] = std::max(res, (R)0);
  }
  (void)M;
}

inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }

}  // namespace

namespace {

// ---------------------------------------------------------------------------------------------------
// forward.cu:155-256 preprocessCUDA (with auxiliary.h:139-164 in_frustum).  Null pointers select code
// paths exactly as in the reference (cov3D_precomp / colors_precomp / scales).
template <typename R>
void preprocess_fwd(int P, int D, int M, const R* means, const R* scales, R mod, const R* rots, const R* opac,
                    const R* shs, const R* cov3D_precomp, const R* colors_precomp, const R* view, const R* proj,
                    const R* campos, int W, int H, R tanx, R tany, int32_t* radii, R* xy, R* depths, R* cov3Ds,
                    R* rgb, R* conic_op, uint32_t* tiles, uint8_t* clamped) {
  const R fy = H / (2 * tany), fx = W / (2 * tanx);  // rasterizer_impl.cu:226-227
  const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; ++i) {
    radii[i] = 0; tiles[i] = 0;
    V3<R> p = {means[3 * i], means[3 * i + 1], means[3 * i + 2]};
    V3<R> pv = xform4x3(p, view);
    if (pv.z <= (R)0.2f) continue;                       // auxiliary.h:154
    V3<R> ph = xform4x3(p, proj);
    R pw = 1 / (xform_w(p, proj) + (R)0.0000001f);       // forward.cu:199
    R ppx = ph.x * pw, ppy = ph.y * pw;
    const R* cov;
    if (cov3D_precomp) cov = cov3D_precomp + 6 * i;
    else { cov3d_from_scale_rot(scales + 3 * i, mod, rots + 4 * i, cov3Ds + 6 * i); cov = cov3Ds + 6 * i; }
    Mat3<R> T, Wm; V3<R> t; R xm, ym;
    ewa_T(p, fx, fy, tanx, tany, view, T, Wm, t, xm, ym);
    Mat3<R> c2 = mul(mul(transpose(T), transpose(sym3(cov))), T);  // forward.cu:106
    R a = c2.m[0][0] + (R)0.3f, b = c2.m[0][1], c = c2.m[1][1] + (R)0.3f;
    R det = a * c - b * b;
    if (det == 0) continue;
    R dinv = 1 / det;
    R con[3] = {c * dinv, -b * dinv, a * dinv};
    R mid = (R)0.5 * (a + c);
    R l1 = mid + std::sqrt(std::max((R)0.1f, mid * mid - det));
    R l2 = mid - std::sqrt(std::max((R)0.1f, mid * mid - det));
    R rad = std::ceil(3 * std::sqrt(std::max(l1, l2)));
    R px = ndc2pix(ppx, W), py = ndc2pix(ppy, H);
    int x0, y0, x1, y1;
    get_rect(px, py, (int)rad, gx, gy, x0, y0, x1, y1);
    if ((x1 - x0) * (y1 - y0) == 0) continue;
    if (!colors_precomp) sh_to_rgb(D, M, p, campos, shs + (size_t)3 * M * i, rgb + 3 * i, clamped + 3 * i);
    depths[i] = pv.z; radii[i] = (int)rad; xy[2 * i] = px; xy[2 * i + 1] = py;
    conic_op[4 * i] = con[0]; conic_op[4 * i + 1] = con[1]; conic_op[4 * i + 2] = con[2];
    conic_op[4 * i + 3] = opac[i];
    tiles[i] = (uint32_t)((y1 - y0) * (x1 - x0));
  }
}

// rasterizer_impl.cu:35-50
uint32_t higher_msb(uint32_t n) {
  uint32_t msb = sizeof(n) * 4, step = msb;
  while (step > 1) { step /= 2; if (n >> msb) msb += step; else msb -= step; }
  if (n >> msb) msb++;
  return msb;
}

// rasterizer_impl.cu:70-111 (duplicateWithKeys), :307-312 (stable radix sort on bits [0,32+bit)),
// :116-138 (identifyTileRanges).  keys/vals must hold R = sum(tiles) entries.
template <typename R>
void binning(int P, int W, int H, const R* xy, const R* depths, const int32_t* radii, const uint32_t* tiles,
             uint64_t* keys_sorted, uint32_t* vals_sorted, uint32_t* ranges /* 2*T */) {
  const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
  size_t Rn = 0;
  for (int i = 0; i < P; ++i) Rn += tiles[i];
  std::vector<uint64_t> keys(Rn);
  std::vector<uint32_t> vals(Rn);
  size_t off = 0;
  for (int i = 0; i < P; ++i) {
    if (radii[i] <= 0) continue;
    int x0, y0, x1, y1;
    get_rect(xy[2 * i], xy[2 * i + 1], radii[i], gx, gy, x0, y0, x1, y1);
    uint32_t dbits = f2u((float)depths[i]);
    for (int y = y0; y < y1; ++y)
      for (int x = x0; x < x1; ++x) {
        keys[off] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | dbits;
        vals[off] = (uint32_t)i;
        ++off;
      }
  }
  const int bit = (int)higher_msb((uint32_t)(gx * gy));
  const uint64_t mask = (bit + 32 >= 64) ? ~0ull : ((1ull << (32 + bit)) - 1);
  std::vector<uint32_t> perm(Rn);
  std::iota(perm.begin(), perm.end(), 0u);
  std::stable_sort(perm.begin(), perm.end(), [&](uint32_t a, uint32_t b) { return (keys[a] & mask) < (keys[b] & mask); });
  for (size_t i = 0; i < Rn; ++i) { keys_sorted[i] = keys[perm[i]]; vals_sorted[i] = vals[perm[i]]; }
  std::memset(ranges, 0, sizeof(uint32_t) * 2 * gx * gy);
  for (size_t i = 0; i < Rn; ++i) {
    uint32_t cur = (uint32_t)(keys_sorted[i] >> 32);
    if (i == 0) ranges[2 * cur] = 0;
    else {
      uint32_t prev = (uint32_t)(keys_sorted[i - 1] >> 32);
      if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; }
    }
    if (i == Rn - 1) ranges[2 * cur + 1] = (uint32_t)Rn;
  }
}

template <typename R> inline R real_exp(R x);
template <> inline float real_exp<float>(float x) { return expf(x); }
template <> inline double real_exp<double>(double x) { return std::exp(x); }

// forward.cu:261-395 renderCUDA.  One pixel at a time; the 256-wide staging rounds of the reference do not
// change any per-pixel result (a pixel stops at its own `done`), so they are not reproduced.
template <typename R>
void render_fwd(int W, int H, int F, const uint32_t* ranges, const uint32_t* point_list, const R* xy,
                const R* feat, const R* lang, const R* depths, const R* conic_op, const R* bg, int include_feature,
                R* final_T, uint32_t* n_contrib, R* out_color, R* out_lang, R* out_depth) {
  const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
  const size_t HW = (size_t)W * H;
#pragma omp parallel for schedule(dynamic, 1)
  for (int tile = 0; tile < gx * gy; ++tile) {
    const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
    const int tx = tile % gx, ty = tile / gx;
    for (int ly = 0; ly < TILE; ++ly)
      for (int lx = 0; lx < TILE; ++lx) {
        const int pxi = tx * TILE + lx, pyi = ty * TILE + ly;
        if (pxi >= W || pyi >= H) continue;
        const R pfx = (R)pxi, pfy = (R)pyi;
        R T = 1, C[3] = {0, 0, 0}, Dp = 0, Fa[16] = {0};
        uint32_t contributor = 0, last = 0;
        for (uint32_t k = r0; k < r1; ++k) {
          ++contributor;
          const uint32_t id = point_list[k];
          const R dx = xy[2 * id] - pfx, dy = xy[2 * id + 1] - pfy;
          const R* co = conic_op + 4 * id;
          const R power = (R)-0.5 * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
          if (power > 0) continue;
          const R alpha = std::min((R)0.99f, co[3] * real_exp<R>(power));
          if (alpha < (R)(1.0f / 255.0f)) continue;
          const R test_T = T * (1 - alpha);
          if (test_T < (R)0.0001f) break;
          const R w = alpha * T;
          for (int ch = 0; ch < 3; ++ch) C[ch] += feat[3 * id + ch] * w;
          if (include_feature) for (int ch = 0; ch < F; ++ch) Fa[ch] += lang[(size_t)F * id + ch] * w;
          Dp += depths[id] * w;
          T = test_T;
          last = contributor;
        }
        const size_t pid = (size_t)W * pyi + pxi;
        final_T[pid] = T; n_contrib[pid] = last;
        for (int ch = 0; ch < 3; ++ch) out_color[ch * HW + pid] = C[ch] + T * bg[ch];
        if (include_feature) for (int ch = 0; ch < F; ++ch) out_lang[ch * HW + pid] = Fa[ch];
        out_depth[pid] = Dp;
      }
  }
}

}  // namespace

namespace {

inline void acc(double* p, double v) {
#pragma omp atomic
  *p += v;
}

// backward.cu:434-674 renderCUDA (backward).  Per-pixel recurrences are restated exactly; the reference's
// global float atomics become double-precision sums (order-independent to ~1e-16), so this is the
// "infinitely careful" value the fp32 atomics of both GPU implementations scatter around.
// Outputs (double accumulators, P-sized, zero-initialised by the caller):
//   g_mean2D[2P] g_conic[3P] (.x,.y,.w of the float4) g_opacity[P] g_color[3P] g_lang[F*P] g_depth[P]
template <typename R>
void render_bwd(int W, int H, int F, const uint32_t* ranges, const uint32_t* point_list, const R* bg, const R* xy,
                const R* conic_op, const R* colors, const R* lang, const R* depths, const R* final_T,
                const uint32_t* n_contrib, const R* dpix, const R* dpix_F, const R* dpix_depth, int include_feature,
                double* g_mean2D, double* g_conic, double* g_opacity, double* g_color, double* g_lang,
                double* g_depth) {
  const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
  const size_t HW = (size_t)W * H;
  const R ddelx_dx = (R)(0.5 * W), ddely_dy = (R)(0.5 * H);  // backward.cu:539-540
#pragma omp parallel for schedule(dynamic, 1)
  for (int tile = 0; tile < gx * gy; ++tile) {
    const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
    const int tx = tile % gx, ty = tile / gx;
    for (int ly = 0; ly < TILE; ++ly)
      for (int lx = 0; lx < TILE; ++lx) {
        const int pxi = tx * TILE + lx, pyi = ty * TILE + ly;
        if (pxi >= W || pyi >= H) continue;
        const size_t pid = (size_t)W * pyi + pxi;
        const R pfx = (R)pxi, pfy = (R)pyi;
        const R T_final = final_T[pid];
        R T = T_final;
        const uint32_t last_contributor = n_contrib[pid];
        R arec[3] = {0, 0, 0}, lastc[3] = {0, 0, 0}, dp[3];
        R arecF[16] = {0}, lastF[16] = {0}, dpF[16] = {0};
        R arecD = 0, lastD = 0, last_alpha = 0;
        for (int ch = 0; ch < 3; ++ch) dp[ch] = dpix[ch * HW + pid];
        const R dpD = dpix_depth[pid];
        if (include_feature) for (int ch = 0; ch < F; ++ch) dpF[ch] = dpix_F[ch * HW + pid];
        for (uint32_t k = r1; k-- > r0;) {
          if (k - r0 >= last_contributor) continue;       // backward.cu:578-580
          const uint32_t id = point_list[k];
          const R dx = xy[2 * id] - pfx, dy = xy[2 * id + 1] - pfy;
          const R* co = conic_op + 4 * id;
          const R power = (R)-0.5 * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
          if (power > 0) continue;
          const R G = real_exp<R>(power);
          const R alpha = std::min((R)0.99f, co[3] * G);
          if (alpha < (R)(1.0f / 255.0f)) continue;
          T = T / (1 - alpha);                             // :595
          const R w = alpha * T;
          R dL_dalpha = 0;
          for (int ch = 0; ch < 3; ++ch) {
            const R c = colors[3 * id + ch];
            arec[ch] = last_alpha * lastc[ch] + (1 - last_alpha) * arec[ch];
            lastc[ch] = c;
            dL_dalpha += (c - arec[ch]) * dp[ch];
            acc(&g_color[3 * id + ch], (double)(w * dp[ch]));
          }
          if (include_feature)
            for (int ch = 0; ch < F; ++ch) {
              const R f = lang[(size_t)F * id + ch];
              arecF[ch] = last_alpha * lastF[ch] + (1 - last_alpha) * arecF[ch];
              lastF[ch] = f;
              dL_dalpha += (f - arecF[ch]) * dpF[ch];
              acc(&g_lang[(size_t)F * id + ch], (double)(w * dpF[ch]));
            }
          const R cd = depths[id];
          arecD = last_alpha * lastD + (1 - last_alpha) * arecD;
          lastD = cd;
          dL_dalpha += (cd - arecD) * dpD;
          acc(&g_depth[id], (double)(w * dpD));
          dL_dalpha *= T;
          last_alpha = alpha;
          R bgdot = 0;
          for (int ch = 0; ch < 3; ++ch) bgdot += bg[ch] * dp[ch];
          dL_dalpha += (-T_final / (1 - alpha)) * bgdot;   // :648-651
          const R dL_dG = co[3] * dL_dalpha;
          const R gdx = G * dx, gdy = G * dy;
          const R dG_ddelx = -gdx * co[0] - gdy * co[1];
          const R dG_ddely = -gdy * co[2] - gdx * co[1];
          acc(&g_mean2D[2 * id], (double)(dL_dG * dG_ddelx * ddelx_dx));
          acc(&g_mean2D[2 * id + 1], (double)(dL_dG * dG_ddely * ddely_dy));
          acc(&g_conic[3 * id], (double)((R)-0.5 * gdx * dx * dL_dG));
          acc(&g_conic[3 * id + 1], (double)((R)-0.5 * gdx * dy * dL_dG));
          acc(&g_conic[3 * id + 2], (double)((R)-0.5 * gdy * dy * dL_dG));
          acc(&g_opacity[id], (double)(G * dL_dalpha));
        }
      }
  }
}

// backward.cu:144-281 computeCov2DCUDA.  dL_dconic is (x, y, w) of the reference's float4 view.
template <typename R>
void cov2d_bwd(int P, const R* means, const int32_t* radii, const R* cov3Ds, R fx, R fy, R tanx, R tany,
               const R* view, const R* dL_dconic3, R* dL_dmeans, R* dL_dcov) {
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; ++i) {
    if (!(radii[i] > 0)) continue;
    const R* c3 = cov3Ds + 6 * i;
    V3<R> mean = {means[3 * i], means[3 * i + 1], means[3 * i + 2]};
    const R dcx = dL_dconic3[3 * i], dcy = dL_dconic3[3 * i + 1], dcz = dL_dconic3[3 * i + 2];
    Mat3<R> T, Wm; V3<R> t; R xm, ym;
    ewa_T(mean, fx, fy, tanx, tany, view, T, Wm, t, xm, ym);
    Mat3<R> V = sym3(c3);
    Mat3<R> c2 = mul(mul(transpose(T), transpose(V)), T);
    const R a = c2.m[0][0] + (R)0.3f, b = c2.m[0][1], c = c2.m[1][1] + (R)0.3f;
    const R denom = a * c - b * b;
    R da = 0, db = 0, dc = 0;
    const R d2i = 1 / ((denom * denom) + (R)0.0000001f);
    R* o = dL_dcov + 6 * i;
    auto Tm = [&](int ci, int ri) { return T.m[ci][ri]; };
    if (d2i != 0) {
      da = d2i * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
      dc = d2i * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
      db = d2i * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
      o[0] = Tm(0, 0) * Tm(0, 0) * da + Tm(0, 0) * Tm(1, 0) * db + Tm(1, 0) * Tm(1, 0) * dc;
      o[3] = Tm(0, 1) * Tm(0, 1) * da + Tm(0, 1) * Tm(1, 1) * db + Tm(1, 1) * Tm(1, 1) * dc;
      o[5] = Tm(0, 2) * Tm(0, 2) * da + Tm(0, 2) * Tm(1, 2) * db + Tm(1, 2) * Tm(1, 2) * dc;
      o[1] = 2 * Tm(0, 0) * Tm(0, 1) * da + (Tm(0, 0) * Tm(1, 1) + Tm(0, 1) * Tm(1, 0)) * db + 2 * Tm(1, 0) * Tm(1, 1) * dc;
      o[2] = 2 * Tm(0, 0) * Tm(0, 2) * da + (Tm(0, 0) * Tm(1, 2) + Tm(0, 2) * Tm(1, 0)) * db + 2 * Tm(1, 0) * Tm(1, 2) * dc;
      o[4] = 2 * Tm(0, 2) * Tm(0, 1) * da + (Tm(0, 1) * Tm(1, 2) + Tm(0, 2) * Tm(1, 1)) * db + 2 * Tm(1, 1) * Tm(1, 2) * dc;
    } else {
      for (int k = 0; k < 6; ++k) o[k] = 0;
    }
    // backward.cu:238-250: dL/dT (upper 2x3).  rowdot(r, k) = sum_j T[r][j] * Vrk[k][j]
    auto rowdot = [&](int r, int k) { return Tm(r, 0) * V.m[k][0] + Tm(r, 1) * V.m[k][1] + Tm(r, 2) * V.m[k][2]; };
    const R dT00 = 2 * rowdot(0, 0) * da + rowdot(1, 0) * db;
    const R dT01 = 2 * rowdot(0, 1) * da + rowdot(1, 1) * db;
    const R dT02 = 2 * rowdot(0, 2) * da + rowdot(1, 2) * db;
    const R dT10 = 2 * rowdot(1, 0) * dc + rowdot(0, 0) * db;
    const R dT11 = 2 * rowdot(1, 1) * dc + rowdot(0, 1) * db;
    const R dT12 = 2 * rowdot(1, 2) * dc + rowdot(0, 2) * db;
    // :254-257
    const R dJ00 = Wm.m[0][0] * dT00 + Wm.m[0][1] * dT01 + Wm.m[0][2] * dT02;
    const R dJ02 = Wm.m[2][0] * dT00 + Wm.m[2][1] * dT01 + Wm.m[2][2] * dT02;
    const R dJ11 = Wm.m[1][0] * dT10 + Wm.m[1][1] * dT11 + Wm.m[1][2] * dT12;
    const R dJ12 = Wm.m[2][0] * dT10 + Wm.m[2][1] * dT11 + Wm.m[2][2] * dT12;
    const R tz = 1 / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
    const R dtx = xm * -fx * tz2 * dJ02;
    const R dty = ym * -fy * tz2 * dJ12;
    const R dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2 * fx * t.x) * tz3 * dJ02 + (2 * fy * t.y) * tz3 * dJ12;
    // auxiliary.h:89-97 transformVec4x3Transpose; assignment (not +=), backward.cu:280
    dL_dmeans[3 * i] = view[0] * dtx + view[1] * dty + view[2] * dtz;
    dL_dmeans[3 * i + 1] = view[4] * dtx + view[5] * dty + view[6] * dtz;
    dL_dmeans[3 * i + 2] = view[8] * dtx + view[9] * dty + view[10] * dtz;
  }
}

}  // namespace

namespace {

// backward.cu:20-139 SH backward: writes dL_dsh (un-clamped channels), returns d/dmean contribution.
template <typename R>
void sh_bwd(int deg, int M, const V3<R>& pos, const R* campos, const R* sh, const uint8_t* clamped, const R* dL_dcolor,
            R* dL_dsh, R dmean[3]) {
  const R ox = pos.x - campos[0], oy = pos.y - campos[1], oz = pos.z - campos[2];
  const R len = std::sqrt(ox * ox + oy * oy + oz * oz);
  const R x = ox / len, y = oy / len, z = oz / len;
  R dRGB[3];
  for (int c = 0; c < 3; ++c) dRGB[c] = dL_dcolor[c] * (clamped[c] ? 0 : 1);
  R dx3[3] = {0, 0, 0}, dy3[3] = {0, 0, 0}, dz3[3] = {0, 0, 0};
  auto S = [&](int k, int c) { return sh[3 * k + c]; };
  auto put = [&](int k, R coef) { for (int c = 0; c < 3; ++c) dL_dsh[3 * k + c] = coef * dRGB[c]; };
  put(0, (R)C0);
  if (deg > 0) {
    put(1, -(R)C1 * y); put(2, (R)C1 * z); put(3, -(R)C1 * x);
    for (int c = 0; c < 3; ++c) { dx3[c] = -(R)C1 * S(3, c); dy3[c] = -(R)C1 * S(1, c); dz3[c] = (R)C1 * S(2, c); }
    if (deg > 1) {
      const R xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      put(4, (R)C2[0] * xy); put(5, (R)C2[1] * yz); put(6, (R)C2[2] * (2 * zz - xx - yy));
      put(7, (R)C2[3] * xz); put(8, (R)C2[4] * (xx - yy));
      for (int c = 0; c < 3; ++c) {
        dx3[c] += (R)C2[0] * y * S(4, c) + (R)C2[2] * 2 * -x * S(6, c) + (R)C2[3] * z * S(7, c) + (R)C2[4] * 2 * x * S(8, c);
        dy3[c] += (R)C2[0] * x * S(4, c) + (R)C2[1] * z * S(5, c) + (R)C2[2] * 2 * -y * S(6, c) + (R)C2[4] * 2 * -y * S(8, c);
        dz3[c] += (R)C2[1] * y * S(5, c) + (R)C2[2] * 2 * 2 * z * S(6, c) + (R)C2[3] * x * S(7, c);
      }
      if (deg > 2) {
        put(9, (R)C3[0] * y * (3 * xx - yy)); put(10, (R)C3[1] * xy * z); put(11, (R)C3[2] * y * (4 * zz - xx - yy));
        put(12, (R)C3[3] * z * (2 * zz - 3 * xx - 3 * yy)); put(13, (R)C3[4] * x * (4 * zz - xx - yy));
        put(14, (R)C3[5] * z * (xx - yy)); put(15, (R)C3[6] * x * (xx - 3 * yy));
        for (int c = 0; c < 3; ++c) {
          dx3[c] += (R)C3[0] * S(9, c) * 3 * 2 * xy + (R)C3[1] * S(10, c) * yz + (R)C3[2] * S(11, c) * -2 * xy +
                    (R)C3[3] * S(12, c) * -3 * 2 * xz + (R)C3[4] * S(13, c) * (-3 * xx + 4 * zz - yy) +
                    (R)C3[5] * S(14, c) * 2 * xz + (R)C3[6] * S(15, c) * 3 * (xx - yy);
          dy3[c] += (R)C3[0] * S(9, c) * 3 * (xx - yy) + (R)C3[1] * S(10, c) * xz +
                    (R)C3[2] * S(11, c) * (-3 * yy + 4 * zz - xx) + (R)C3[3] * S(12, c) * -3 * 2 * yz +
                    (R)C3[4] * S(13, c) * -2 * xy + (R)C3[5] * S(14, c) * -2 * yz + (R)C3[6] * S(15, c) * -3 * 2 * xy;
          dz3[c] += (R)C3[1] * S(10, c) * xy + (R)C3[2] * S(11, c) * 4 * 2 * yz +
                    (R)C3[3] * S(12, c) * 3 * (2 * zz - xx - yy) + (R)C3[4] * S(13, c) * 4 * 2 * xz +
                    (R)C3[5] * S(14, c) * (xx - yy);
        }
      }
    }
  }
  const R ddx = dx3[0] * dRGB[0] + dx3[1] * dRGB[1] + dx3[2] * dRGB[2];
  const R ddy = dy3[0] * dRGB[0] + dy3[1] * dRGB[1] + dy3[2] * dRGB[2];
  const R ddz = dz3[0] * dRGB[0] + dz3[1] * dRGB[1] + dz3[2] * dRGB[2];
  // auxiliary.h:107-117 dnormvdv
  const R s2 = ox * ox + oy * oy + oz * oz;
  const R inv = 1 / std::sqrt(s2 * s2 * s2);
  dmean[0] = ((s2 - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * inv;
  dmean[1] = (-ox * oy * ddx + (s2 - oy * oy) * ddy - oz * oy * ddz) * inv;
  dmean[2] = (-ox * oz * ddx - oy * oz * ddy + (s2 - oz * oz) * ddz) * inv;
  (void)M;
}

// backward.cu:285-348: cov3D -> scale / rotation gradients (no quaternion-normalisation Jacobian, :347)
template <typename R>
void cov3d_bwd(const R* s, R mod, const R* q, const R* g /*6*/, R* dscale, R* drot) {
  Mat3<R> M, Rm; R sc[3];
  rot_scale_M(s, mod, q, M, Rm, sc);
  const R r = q[0], x = q[1], y = q[2], z = q[3];
  Mat3<R> dSig;
  dSig.m[0][0] = g[0]; dSig.m[0][1] = (R)0.5 * g[1]; dSig.m[0][2] = (R)0.5 * g[2];
  dSig.m[1][0] = (R)0.5 * g[1]; dSig.m[1][1] = g[3]; dSig.m[1][2] = (R)0.5 * g[4];
  dSig.m[2][0] = (R)0.5 * g[2]; dSig.m[2][1] = (R)0.5 * g[4]; dSig.m[2][2] = g[5];
  Mat3<R> M2 = M;
  for (int c = 0; c < 3; ++c) for (int rr = 0; rr < 3; ++rr) M2.m[c][rr] *= 2;
  Mat3<R> dM = mul(M2, dSig);
  Mat3<R> Rt = transpose(Rm), dMt = transpose(dM);
  for (int k = 0; k < 3; ++k)
    dscale[k] = Rt.m[k][0] * dMt.m[k][0] + Rt.m[k][1] * dMt.m[k][1] + Rt.m[k][2] * dMt.m[k][2];
  for (int k = 0; k < 3; ++k) for (int rr = 0; rr < 3; ++rr) dMt.m[k][rr] *= sc[k];
  auto D = [&](int c, int rr) { return dMt.m[c][rr]; };
  drot[0] = 2 * z * (D(0, 1) - D(1, 0)) + 2 * y * (D(2, 0) - D(0, 2)) + 2 * x * (D(1, 2) - D(2, 1));
  drot[1] = 2 * y * (D(1, 0) + D(0, 1)) + 2 * z * (D(2, 0) + D(0, 2)) + 2 * r * (D(1, 2) - D(2, 1)) - 4 * x * (D(2, 2) + D(1, 1));
  drot[2] = 2 * x * (D(1, 0) + D(0, 1)) + 2 * r * (D(2, 0) - D(0, 2)) + 2 * z * (D(1, 2) + D(2, 1)) - 4 * y * (D(2, 2) + D(0, 0));
  drot[3] = 2 * r * (D(0, 1) - D(1, 0)) + 2 * x * (D(2, 0) + D(0, 2)) + 2 * y * (D(1, 2) + D(2, 1)) - 4 * z * (D(1, 1) + D(0, 0));
}

// backward.cu:353-431 preprocessCUDA (backward).  dL_dmeans already holds the cov2D part (cov2d_bwd).
template <typename R>
void preprocess_bwd(int P, int D, int M, const R* means, const int32_t* radii, const R* shs, const uint8_t* clamped,
                    const R* scales, const R* rots, R mod, const R* view, const R* proj, const R* campos,
                    const R* dL_dmean2D /*3P, .z unused*/, R* dL_dmeans, const R* dL_dcolor, const R* dL_ddepth,
                    const R* dL_dcov3D, R* dL_dsh, R* dL_dscale, R* dL_drot) {
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; ++i) {
    if (!(radii[i] > 0)) continue;
    V3<R> m = {means[3 * i], means[3 * i + 1], means[3 * i + 2]};
    const R mw = 1 / (xform_w(m, proj) + (R)0.0000001f);
    const R mul1 = (proj[0] * m.x + proj[4] * m.y + proj[8] * m.z + proj[12]) * mw * mw;
    const R mul2 = (proj[1] * m.x + proj[5] * m.y + proj[9] * m.z + proj[13]) * mw * mw;
    const R gx = dL_dmean2D[3 * i], gy = dL_dmean2D[3 * i + 1];
    R d[3];
    d[0] = (proj[0] * mw - proj[3] * mul1) * gx + (proj[1] * mw - proj[3] * mul2) * gy;
    d[1] = (proj[4] * mw - proj[7] * mul1) * gx + (proj[5] * mw - proj[7] * mul2) * gy;
    d[2] = (proj[8] * mw - proj[11] * mul1) * gx + (proj[9] * mw - proj[11] * mul2) * gy;
    for (int k = 0; k < 3; ++k) dL_dmeans[3 * i + k] += d[k];
    const R mul3 = view[2] * m.x + view[6] * m.y + view[10] * m.z + view[14];  // :411
    const R gd = dL_ddepth[i];
    dL_dmeans[3 * i] += (view[2] - view[3] * mul3) * gd;
    dL_dmeans[3 * i + 1] += (view[6] - view[7] * mul3) * gd;
    dL_dmeans[3 * i + 2] += (view[10] - view[11] * mul3) * gd;
    if (shs) {
      R dm[3];
      sh_bwd(D, M, m, campos, shs + (size_t)3 * M * i, clamped + 3 * i, dL_dcolor + 3 * i, dL_dsh + (size_t)3 * M * i, dm);
      for (int k = 0; k < 3; ++k) dL_dmeans[3 * i + k] += dm[k];
    }
    if (scales) cov3d_bwd(scales + 3 * i, mod, rots + 4 * i, dL_dcov3D + 6 * i, dL_dscale + 3 * i, dL_drot + 4 * i);
  }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------
// C ABI (ctypes).  Every entry point exists as <name>_f32 and <name>_f64.
#define ORC_EXPORTS(SUF, R)                                                                                          \
  ORC_API void orc_preprocess_fwd_##SUF(int P, int D, int M, const R* means, const R* scales, R mod, const R* rots,  \
                                        const R* opac, const R* shs, const R* cov3D_precomp, const R* colors_precomp, \
                                        const R* view, const R* proj, const R* campos, int W, int H, R tanx, R tany,  \
                                        int32_t* radii, R* xy, R* depths, R* cov3Ds, R* rgb, R* conic_op,             \
                                        uint32_t* tiles, uint8_t* clamped) {                                          \
    preprocess_fwd<R>(P, D, M, means, scales, mod, rots, opac, shs, cov3D_precomp, colors_precomp, view, proj,       \
                      campos, W, H, tanx, tany, radii, xy, depths, cov3Ds, rgb, conic_op, tiles, clamped);            \
  }                                                                                                                  \
  ORC_API void orc_binning_##SUF(int P, int W, int H, const R* xy, const R* depths, const int32_t* radii,            \
                                 const uint32_t* tiles, uint64_t* keys_sorted, uint32_t* vals_sorted,                 \
                                 uint32_t* ranges) {                                                                  \
    binning<R>(P, W, H, xy, depths, radii, tiles, keys_sorted, vals_sorted, ranges);                                 \
  }                                                                                                                  \
  ORC_API void orc_render_fwd_##SUF(int W, int H, int F, const uint32_t* ranges, const uint32_t* point_list,         \
                                    const R* xy, const R* feat, const R* lang, const R* depths, const R* conic_op,    \
                                    const R* bg, int include_feature, R* final_T, uint32_t* n_contrib, R* out_color,  \
                                    R* out_lang, R* out_depth) {                                                      \
    render_fwd<R>(W, H, F, ranges, point_list, xy, feat, lang, depths, conic_op, bg, include_feature, final_T,       \
                  n_contrib, out_color, out_lang, out_depth);                                                        \
  }                                                                                                                  \
  ORC_API void orc_render_bwd_##SUF(int W, int H, int F, const uint32_t* ranges, const uint32_t* point_list,         \
                                    const R* bg, const R* xy, const R* conic_op, const R* colors, const R* lang,      \
                                    const R* depths, const R* final_T, const uint32_t* n_contrib, const R* dpix,      \
                                    const R* dpix_F, const R* dpix_depth, int include_feature, double* g_mean2D,      \
                                    double* g_conic, double* g_opacity, double* g_color, double* g_lang,              \
                                    double* g_depth) {                                                                \
    render_bwd<R>(W, H, F, ranges, point_list, bg, xy, conic_op, colors, lang, depths, final_T, n_contrib, dpix,     \
                  dpix_F, dpix_depth, include_feature, g_mean2D, g_conic, g_opacity, g_color, g_lang, g_depth);      \
  }                                                                                                                  \
  ORC_API void orc_cov2d_bwd_##SUF(int P, const R* means, const int32_t* radii, const R* cov3Ds, int W, int H,       \
                                   R tanx, R tany, const R* view, const R* dL_dconic3, R* dL_dmeans, R* dL_dcov) {   \
    const R fy = H / (2 * tany), fx = W / (2 * tanx);                                                                \
    cov2d_bwd<R>(P, means, radii, cov3Ds, fx, fy, tanx, tany, view, dL_dconic3, dL_dmeans, dL_dcov);                 \
  }                                                                                                                  \
  ORC_API void orc_preprocess_bwd_##SUF(int P, int D, int M, const R* means, const int32_t* radii, const R* shs,     \
                                        const uint8_t* clamped, const R* scales, const R* rots, R mod, const R* view, \
                                        const R* proj, const R* campos, const R* dL_dmean2D, R* dL_dmeans,            \
                                        const R* dL_dcolor, const R* dL_ddepth, const R* dL_dcov3D, R* dL_dsh,        \
                                        R* dL_dscale, R* dL_drot) {                                                   \
    preprocess_bwd<R>(P, D, M, means, radii, shs, clamped, scales, rots, mod, view, proj, campos, dL_dmean2D,        \
                      dL_dmeans, dL_dcolor, dL_ddepth, dL_dcov3D, dL_dsh, dL_dscale, dL_drot);                       \
  }

ORC_EXPORTS(f32, float)
ORC_EXPORTS(f64, double)

// rasterizer_impl.cu:54-66 checkFrustum / auxiliary.h:139-164
ORC_API void orc_mark_visible_f32(int P, const float* means, const float* view, uint8_t* present) {
  for (int i = 0; i < P; ++i) {
    V3<float> p = {means[3 * i], means[3 * i + 1], means[3 * i + 2]};
    present[i] = xform4x3(p, view).z > 0.2f;
  }
}

ORC_API uint32_t orc_higher_msb(uint32_t n) { return higher_msb(n); }
