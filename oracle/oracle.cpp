// oracle.cpp — CPU restatement of the reference rasterizer math.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path: only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library.
//
// PARITY STATUS: the reference ships no tests, golden vectors or CPU rasterizer for this path
// (SURVEY.md §4), so this restatement is pinned by (a) the unmodified reference CUDA extension built
// into baseline/_ref and run on the GPU box by tests/test_vs_reference.py (bit-exact radii / keys /
// sorted list / ranges, 1e-5 images, 1e-4 gradients), (b) fixtures under tests/golden/ generated from
// the reference's own Python re-statements of SH->RGB and cov3D (utils/sh_utils.py,
// utils/general_utils.py), and (c) a float64 finite-difference gradcheck of this file.
//
// Every function cites the reference file:line it follows.  RAST = submodules/langsplat-rasterization.
// The float instantiation mirrors nvcc's FMA contraction of the reference source:
// for `p1 + p2 + p3 (+c)` nvcc emits  t = a2*b2; t = fma(a1,b1,t); t = fma(a3,b3,t); t = t + c.
// Build with -ffp-contract=off so gcc adds no contractions of its own.
//
// Templated on the scalar type so the same code runs in float (parity) and double (gradcheck).

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <numeric>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

constexpr int BLOCK_X = 16, BLOCK_Y = 16;  // RAST/cuda_rasterizer/config.h:17-18

// RAST/cuda_rasterizer/auxiliary.h:22-39
constexpr double SH_C0 = 0.28209479177387814;
constexpr double SH_C1 = 0.4886025119029199;
constexpr double SH_C2[5] = {1.0925484305920792, -1.0925484305920792, 0.31539156525252005,
                             -1.0925484305920792, 0.5462742152960396};
constexpr double SH_C3[7] = {-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
                             -0.4570457994644658, 1.445305721320277, -0.5900435899266435};

template <class R>
inline R c_(double v) {  // constants are float literals in the reference
  return (R)(float)v;
}

// glm::mat3, column major: m[c][r]
template <class R>
struct M3 {
  R m[3][3];
};

// RAST/third_party/glm/glm/detail/type_mat3x3.inl:486-517 with nvcc's contraction
template <class R>
inline M3<R> mul(const M3<R>& A, const M3<R>& B) {
  M3<R> C;
  for (int c = 0; c < 3; c++)
    for (int r = 0; r < 3; r++) {
      R t = A.m[1][r] * B.m[c][1];
      t = std::fma(A.m[0][r], B.m[c][0], t);
      t = std::fma(A.m[2][r], B.m[c][2], t);
      C.m[c][r] = t;
    }
  return C;
}
template <class R>
inline M3<R> transpose(const M3<R>& A) {
  M3<R> C;
  for (int c = 0; c < 3; c++)
    for (int r = 0; r < 3; r++) C.m[c][r] = A.m[r][c];
  return C;
}

// RAST/cuda_rasterizer/auxiliary.h:58-77 — one output row of transformPoint4x3 / 4x4
template <class R>
inline R tp_row(const R* m, int i, R x, R y, R z) {
  R t = m[4 + i] * y;
  t = std::fma(m[i], x, t);
  t = std::fma(m[8 + i], z, t);
  return t + m[12 + i];
}

inline int f2i_rz(double v) {  // cvt.rzi.s32.f32 saturates; NaN -> 0
  if (std::isnan(v)) return 0;
  if (v >= 2147483647.0) return 2147483647;
  if (v <= -2147483648.0) return (int)-2147483648LL;
  return (int)v;
}

// RAST/cuda_rasterizer/auxiliary.h:41-44 (evaluated in double)
template <class R>
inline R ndc2Pix(R v, int S) {
  return (R)(float)(std::fma((double)v + 1.0, (double)S, -1.0) * 0.5);
}
template <>
inline double ndc2Pix<double>(double v, int S) {
  return ((v + 1.0) * S - 1.0) * 0.5;
}

// RAST/cuda_rasterizer/auxiliary.h:46-56
template <class R>
inline void getRect(R px, R py, int max_radius, int gx, int gy, uint32_t* rmin, uint32_t* rmax) {
  R r = (R)max_radius;
  rmin[0] = (uint32_t)std::min<int64_t>(gx, std::max(0, f2i_rz((px - r) * (R)0.0625)));
  rmin[1] = (uint32_t)std::min<int64_t>(gy, std::max(0, f2i_rz((py - r) * (R)0.0625)));
  rmax[0] = (uint32_t)std::min<int64_t>(gx, std::max(0, f2i_rz((((px + r) + (R)16) - (R)1) * (R)0.0625)));
  rmax[1] = (uint32_t)std::min<int64_t>(gy, std::max(0, f2i_rz((((py + r) + (R)16) - (R)1) * (R)0.0625)));
}

// rotation matrix from the (un-normalised, as given) quaternion — RAST/cuda_rasterizer/forward.cu:127-139
template <class R>
inline M3<R> quatR(const R* q) {
  R r = q[0], x = q[1], y = q[2], z = q[3];
  auto two = [](R v) { return v + v; };
  M3<R> Rm;
  Rm.m[0][0] = (R)1 - two(y * y + z * z);  // both products have other uses: no fma in the reference SASS
  Rm.m[0][1] = two(std::fma(x, y, -(r * z)));
  Rm.m[0][2] = two(std::fma(r, y, x * z));
  Rm.m[1][0] = two(std::fma(x, y, r * z));
  Rm.m[1][1] = (R)1 - two(std::fma(x, x, z * z));
  Rm.m[1][2] = two(std::fma(y, z, -(r * x)));
  Rm.m[2][0] = two(std::fma(-r, y, x * z));
  Rm.m[2][1] = two(std::fma(y, z, r * x));
  Rm.m[2][2] = (R)1 - two(std::fma(x, x, y * y));
  return Rm;
}

// RAST/cuda_rasterizer/forward.cu:119-153
template <class R>
inline void computeCov3D(const R* scale, R mod, const R* rot, R* cov3D) {
  M3<R> Rm = quatR(rot);
  R s[3] = {mod * scale[0], mod * scale[1], mod * scale[2]};
  M3<R> M;  // M = S * R  ->  M[c][r] = s_r * R[c][r]
  for (int c = 0; c < 3; c++)
    for (int r = 0; r < 3; r++) M.m[c][r] = s[r] * Rm.m[c][r];
  M3<R> Sigma = mul(transpose(M), M);
  cov3D[0] = Sigma.m[0][0];
  cov3D[1] = Sigma.m[0][1];
  cov3D[2] = Sigma.m[0][2];
  cov3D[3] = Sigma.m[1][1];
  cov3D[4] = Sigma.m[1][2];
  cov3D[5] = Sigma.m[2][2];
}

template <class R>
struct Cov2DCtx {  // intermediates shared by forward and backward
  R tx, ty, tz, txtz, tytz, limx, limy;
  M3<R> J, W, T, Vrk, cov;
};

// RAST/cuda_rasterizer/forward.cu:75-114 (and the recomputation in backward.cu:155-201)
template <class R>
inline void computeCov2D(const R* mean, R fx, R fy, R tan_fovx, R tan_fovy, const R* cov3D, const R* V,
                         Cov2DCtx<R>& c) {
  R tx = tp_row(V, 0, mean[0], mean[1], mean[2]);
  R ty = tp_row(V, 1, mean[0], mean[1], mean[2]);
  R tz = tp_row(V, 2, mean[0], mean[1], mean[2]);
  c.limx = c_<R>(1.3) * tan_fovx;
  c.limy = c_<R>(1.3) * tan_fovy;
  c.txtz = tx / tz;
  c.tytz = ty / tz;
  tx = std::min(c.limx, std::max(-c.limx, c.txtz)) * tz;
  ty = std::min(c.limy, std::max(-c.limy, c.tytz)) * tz;
  c.tx = tx;
  c.ty = ty;
  c.tz = tz;
  M3<R>& J = c.J;
  std::memset(&J, 0, sizeof(J));
  J.m[0][0] = fx / tz;
  J.m[0][2] = -(fx * tx) / (tz * tz);
  J.m[1][1] = fy / tz;
  J.m[1][2] = -(fy * ty) / (tz * tz);
  M3<R>& W = c.W;
  W.m[0][0] = V[0]; W.m[0][1] = V[4]; W.m[0][2] = V[8];
  W.m[1][0] = V[1]; W.m[1][1] = V[5]; W.m[1][2] = V[9];
  W.m[2][0] = V[2]; W.m[2][1] = V[6]; W.m[2][2] = V[10];
  c.T = mul(W, J);
  M3<R>& K = c.Vrk;
  K.m[0][0] = cov3D[0]; K.m[0][1] = cov3D[1]; K.m[0][2] = cov3D[2];
  K.m[1][0] = cov3D[1]; K.m[1][1] = cov3D[3]; K.m[1][2] = cov3D[4];
  K.m[2][0] = cov3D[2]; K.m[2][1] = cov3D[4]; K.m[2][2] = cov3D[5];
  c.cov = mul(mul(transpose(c.T), transpose(K)), c.T);
  c.cov.m[0][0] += c_<R>(0.3);
  c.cov.m[1][1] += c_<R>(0.3);
}

// RAST/cuda_rasterizer/forward.cu:21-72
template <class R>
inline void shToColor(int deg, int M, const R* mean, const R* campos, const R* sh /*[M][3]*/, R* rgb,
                      uint8_t* clamped) {
  R d[3] = {mean[0] - campos[0], mean[1] - campos[1], mean[2] - campos[2]};
  R len = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
  R x = d[0] / len, y = d[1] / len, z = d[2] / len;
  for (int ch = 0; ch < 3; ch++) {
    auto S = [&](int k) { return sh[k * 3 + ch]; };
    R res = c_<R>(SH_C0) * S(0);
    if (deg > 0) {
      res = res - c_<R>(SH_C1) * y * S(1) + c_<R>(SH_C1) * z * S(2) - c_<R>(SH_C1) * x * S(3);
      if (deg > 1) {
        R xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        res = res + c_<R>(SH_C2[0]) * xy * S(4) + c_<R>(SH_C2[1]) * yz * S(5) +
              c_<R>(SH_C2[2]) * ((R)2 * zz - xx - yy) * S(6) + c_<R>(SH_C2[3]) * xz * S(7) +
              c_<R>(SH_C2[4]) * (xx - yy) * S(8);
        if (deg > 2) {
          res = res + c_<R>(SH_C3[0]) * y * ((R)3 * xx - yy) * S(9) + c_<R>(SH_C3[1]) * xy * z * S(10) +
                c_<R>(SH_C3[2]) * y * ((R)4 * zz - xx - yy) * S(11) +
                c_<R>(SH_C3[3]) * z * ((R)2 * zz - (R)3 * xx - (R)3 * yy) * S(12) +
                c_<R>(SH_C3[4]) * x * ((R)4 * zz - xx - yy) * S(13) + c_<R>(SH_C3[5]) * z * (xx - yy) * S(14) +
                c_<R>(SH_C3[6]) * x * (xx - (R)3 * yy) * S(15);
        }
      }
    }
    res += (R)0.5;
    clamped[ch] = res < 0;
    rgb[ch] = std::max(res, (R)0);
  }
  (void)M;
}

// RAST/cuda_rasterizer/rasterizer_impl.cu:35-50
inline uint32_t getHigherMsb(uint32_t n) {
  uint32_t msb = sizeof(n) * 4, step = msb;
  while (step > 1) {
    step /= 2;
    if (n >> msb) msb += step; else msb -= step;
  }
  if (n >> msb) msb++;
  return msb;
}

template <class R>
inline uint32_t depth_bits(R d) {
  float f = (float)d;
  uint32_t u;
  std::memcpy(&u, &f, 4);
  return u;
}

// ---------------------------------------------------------------------------------------------------
// Stage 1: per-Gaussian preprocess — RAST/cuda_rasterizer/forward.cu:156-257, auxiliary.h:139-164
// ---------------------------------------------------------------------------------------------------
template <class R>
int64_t preprocess_T(int P, int D, int M, const R* means3D, const R* scales, R scale_modifier, const R* rotations,
                     const R* opacities, const R* shs, const R* cov3D_precomp, const R* colors_precomp,
                     const R* viewmatrix, const R* projmatrix, const R* campos, int W, int H, R tan_fovx, R tan_fovy,
                     int32_t* radii, R* means2D, R* depths, R* cov3Ds, R* rgb, R* conic_opacity, uint8_t* clamped,
                     uint32_t* tiles_touched, uint32_t* point_offsets) {
  // RAST/cuda_rasterizer/rasterizer_impl.cu:225-226 (host, f32: int -> float, 2.0f*tan, divide)
  const R focal_y = (R)H / ((R)2 * tan_fovy);
  const R focal_x = (R)W / ((R)2 * tan_fovx);
  const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
#pragma omp parallel for schedule(static)
  for (int idx = 0; idx < P; idx++) {
    radii[idx] = 0;
    tiles_touched[idx] = 0;
    const R* p = means3D + 3 * (size_t)idx;
    // in_frustum
    R hx = tp_row(projmatrix, 0, p[0], p[1], p[2]);
    R hy = tp_row(projmatrix, 1, p[0], p[1], p[2]);
    R hw = tp_row(projmatrix, 3, p[0], p[1], p[2]);
    R p_w = (R)1 / (hw + c_<R>(0.0000001));
    R projx = hx * p_w, projy = hy * p_w;
    R viewz = tp_row(viewmatrix, 2, p[0], p[1], p[2]);
    if (viewz <= c_<R>(0.2)) continue;
    const R* cov3D;
    if (cov3D_precomp) {
      cov3D = cov3D_precomp + 6 * (size_t)idx;
    } else {
      computeCov3D(scales + 3 * (size_t)idx, scale_modifier, rotations + 4 * (size_t)idx, cov3Ds + 6 * (size_t)idx);
      cov3D = cov3Ds + 6 * (size_t)idx;
    }
    Cov2DCtx<R> cc;
    computeCov2D(p, focal_x, focal_y, tan_fovx, tan_fovy, cov3D, viewmatrix, cc);
    R a = cc.cov.m[0][0], b = cc.cov.m[0][1], c = cc.cov.m[1][1];
    R det = std::fma(a, c, -(b * b));
    if (det == (R)0) continue;
    R det_inv = (R)1 / det;
    R conic[3] = {c * det_inv, -b * det_inv, a * det_inv};
    R mid = (R)0.5 * (a + c);
    R s = std::sqrt(std::max(c_<R>(0.1), std::fma(mid, mid, -det)));
    R lambda1 = mid + s, lambda2 = mid - s;
    R my_radius = std::ceil((R)3 * std::sqrt(std::max(lambda1, lambda2)));
    R pix[2] = {ndc2Pix(projx, W), ndc2Pix(projy, H)};
    uint32_t rmin[2], rmax[2];
    int radius_i = f2i_rz((double)my_radius);
    getRect(pix[0], pix[1], radius_i, gx, gy, rmin, rmax);
    if ((rmax[0] - rmin[0]) * (rmax[1] - rmin[1]) == 0) continue;
    if (!colors_precomp)
      shToColor(D, M, p, campos, shs + (size_t)idx * M * 3, rgb + 3 * (size_t)idx, clamped + 3 * (size_t)idx);
    depths[idx] = viewz;
    radii[idx] = radius_i;
    means2D[2 * (size_t)idx] = pix[0];
    means2D[2 * (size_t)idx + 1] = pix[1];
    conic_opacity[4 * (size_t)idx + 0] = conic[0];
    conic_opacity[4 * (size_t)idx + 1] = conic[1];
    conic_opacity[4 * (size_t)idx + 2] = conic[2];
    conic_opacity[4 * (size_t)idx + 3] = opacities[idx];
    tiles_touched[idx] = (rmax[1] - rmin[1]) * (rmax[0] - rmin[0]);
  }
  // RAST/cuda_rasterizer/rasterizer_impl.cu:280 (cub::DeviceScan::InclusiveSum)
  uint32_t acc = 0;
  for (int i = 0; i < P; i++) {
    acc += tiles_touched[i];
    point_offsets[i] = acc;
  }
  return P ? (int64_t)point_offsets[P - 1] : 0;
}

// ---------------------------------------------------------------------------------------------------
// Stage 2: duplicate, sort, ranges — RAST/cuda_rasterizer/rasterizer_impl.cu:70-138, 292-323
// ---------------------------------------------------------------------------------------------------
template <class R>
void binning_T(int P, int W, int H, const int32_t* radii, const R* means2D, const R* depths,
               const uint32_t* point_offsets, int64_t num_rendered, uint64_t* keys_unsorted,
               uint32_t* values_unsorted, uint64_t* keys_sorted, uint32_t* point_list, uint32_t* ranges /*[T][2]*/) {
  const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
  const int T = gx * gy;
#pragma omp parallel for schedule(dynamic, 1024)
  for (int idx = 0; idx < P; idx++) {
    if (radii[idx] > 0) {
      uint32_t off = idx == 0 ? 0 : point_offsets[idx - 1];
      uint32_t rmin[2], rmax[2];
      getRect(means2D[2 * (size_t)idx], means2D[2 * (size_t)idx + 1], radii[idx], gx, gy, rmin, rmax);
      for (uint32_t y = rmin[1]; y < rmax[1]; y++)
        for (uint32_t x = rmin[0]; x < rmax[0]; x++) {
          uint64_t key = (uint64_t)(y * gx + x);
          key <<= 32;
          key |= depth_bits(depths[idx]);
          keys_unsorted[off] = key;
          values_unsorted[off] = (uint32_t)idx;
          off++;
        }
    }
  }
  // Stable ascending sort over bits [0, 32+getHigherMsb(T)) — every higher bit is zero by construction, so a
  // stable sort on the whole key is identical (rasterizer_impl.cu:303-311).  Done as a counting sort on the
  // tile id followed by a stable per-tile sort on the depth bits: same permutation as any stable sort.
  (void)getHigherMsb;
  std::vector<uint32_t> cnt((size_t)T + 1, 0);
  for (int64_t i = 0; i < num_rendered; i++) cnt[(keys_unsorted[i] >> 32) + 1]++;
  for (int t = 0; t < T; t++) cnt[t + 1] += cnt[t];
  {
    std::vector<uint32_t> cur(cnt.begin(), cnt.end() - 1);
    for (int64_t i = 0; i < num_rendered; i++) {
      uint32_t t = (uint32_t)(keys_unsorted[i] >> 32);
      uint32_t d = cur[t]++;
      keys_sorted[d] = keys_unsorted[i];
      point_list[d] = values_unsorted[i];
    }
  }
#pragma omp parallel
  {
    std::vector<std::pair<uint64_t, uint32_t>> tmp;
#pragma omp for schedule(dynamic, 16)
    for (int t = 0; t < T; t++) {
      uint32_t b = cnt[t], e = cnt[t + 1];
      if (e - b < 2) continue;
      tmp.resize(e - b);
      for (uint32_t i = b; i < e; i++) tmp[i - b] = {keys_sorted[i], point_list[i]};
      std::stable_sort(tmp.begin(), tmp.end(), [](const auto& a, const auto& b2) { return a.first < b2.first; });
      for (uint32_t i = b; i < e; i++) {
        keys_sorted[i] = tmp[i - b].first;
        point_list[i] = tmp[i - b].second;
      }
    }
  }
  // identifyTileRanges (rasterizer_impl.cu:116-138) after cudaMemset(ranges, 0) (:313)
  std::memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)T);
  for (int64_t i = 0; i < num_rendered; i++) {
    uint32_t cur = (uint32_t)(keys_sorted[i] >> 32);
    if (i == 0)
      ranges[2 * cur] = 0;
    else {
      uint32_t prev = (uint32_t)(keys_sorted[i - 1] >> 32);
      if (cur != prev) {
        ranges[2 * prev + 1] = (uint32_t)i;
        ranges[2 * cur] = (uint32_t)i;
      }
    }
    if (i == num_rendered - 1) ranges[2 * cur + 1] = (uint32_t)num_rendered;
  }
}

template <class R>
inline R exp_(R x) { return std::exp(x); }
template <>
inline float exp_<float>(float x) { return expf(x); }

// ---------------------------------------------------------------------------------------------------
// Stage 3: tile blend forward — RAST/cuda_rasterizer/forward.cu:262-393
// ---------------------------------------------------------------------------------------------------
template <class R>
void render_T(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const R* means2D, const R* colors,
              const R* lang, const R* conic_opacity, const R* bg, int include_feature, R* final_T,
              uint32_t* n_contrib, R* out_color, R* out_feature, int64_t* k_exam, int64_t* k_contrib) {
  const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
  const size_t N = (size_t)W * H;
  int64_t exam = 0, contrib = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : exam, contrib)
  for (int tile = 0; tile < gx * gy; tile++) {
    const int tx = tile % gx, ty = tile / gx;
    const uint32_t rb = ranges[2 * tile], re = ranges[2 * tile + 1];
    for (int ly = 0; ly < BLOCK_Y; ly++)
      for (int lx = 0; lx < BLOCK_X; lx++) {
        const int px = tx * BLOCK_X + lx, py = ty * BLOCK_Y + ly;
        if (px >= W || py >= H) continue;
        const size_t pix_id = (size_t)W * py + px;
        const R pxf = (R)px, pyf = (R)py;
        R T = 1;
        uint32_t contributor = 0, last_contributor = 0;
        R C[3] = {0, 0, 0}, F[3] = {0, 0, 0};
        for (uint32_t k = rb; k < re; k++) {
          contributor++;
          const uint32_t id = point_list[k];
          const R dx = means2D[2 * (size_t)id] - pxf, dy = means2D[2 * (size_t)id + 1] - pyf;
          const R* co = conic_opacity + 4 * (size_t)id;
          // S = fma(dx, con.x*dx, (con.z*dy)*dy); power = fma(S, -0.5, -((con.y*dx)*dy))
          const R S = std::fma(dx, co[0] * dx, (co[2] * dy) * dy);
          const R power = std::fma(S, (R)-0.5, -((co[1] * dx) * dy));
          if (power > 0) continue;
          const R alpha = std::min(c_<R>(0.99), co[3] * exp_(power));
          if (alpha < (R)(1.0f / 255.0f)) continue;
          const R test_T = T * ((R)1 - alpha);
          if (test_T < c_<R>(0.0001)) break;  // done = true: nothing further is examined
          contrib++;
          for (int ch = 0; ch < 3; ch++) C[ch] += colors[3 * (size_t)id + ch] * alpha * T;
          if (include_feature)
            for (int ch = 0; ch < 3; ch++) F[ch] += lang[3 * (size_t)id + ch] * alpha * T;
          T = test_T;
          last_contributor = contributor;
        }
        exam += last_contributor;
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
        for (int ch = 0; ch < 3; ch++) out_color[ch * N + pix_id] = C[ch] + T * bg[ch];
        if (include_feature)
          for (int ch = 0; ch < 3; ch++) out_feature[ch * N + pix_id] = F[ch];
      }
  }
  if (k_exam) *k_exam = exam;
  if (k_contrib) *k_contrib = contrib;
}

// ---------------------------------------------------------------------------------------------------
// Stage 4: tile blend backward — RAST/cuda_rasterizer/backward.cu:399-593
// Accumulates into dL_dmean2D[P][3], dL_dconic[P][4], dL_dopacity[P], dL_dcolors[P][3], dL_dfeat[P][3]
// (all must be zero on entry, as torch::zeros does in rasterize_points.cu:166-183).
// ---------------------------------------------------------------------------------------------------
template <class R>
void render_bw_T(int P, int W, int H, const uint32_t* ranges, const uint32_t* point_list, const R* bg,
                 const R* means2D, const R* conic_opacity, const R* colors, const R* lang, const R* final_Ts,
                 const uint32_t* n_contrib, const R* dL_dpixels, const R* dL_dpixels_F, int include_feature,
                 R* dL_dmean2D, R* dL_dconic, R* dL_dopacity, R* dL_dcolors, R* dL_dfeat) {
  const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
  const size_t N = (size_t)W * H;
  const R ddelx_dx = (R)0.5 * W, ddely_dy = (R)0.5 * H;
  int nthreads = 1;
#ifdef _OPENMP
  nthreads = omp_get_max_threads();
#endif
  // per-thread private accumulators (14 values per Gaussian), reduced at the end; accumulate in double so the
  // oracle is order-independent to f32 precision (the reference's own order is non-deterministic atomics).
  std::vector<std::vector<double>> acc(nthreads);
#pragma omp parallel
  {
    int tid = 0;
#ifdef _OPENMP
    tid = omp_get_thread_num();
#endif
    std::vector<double>& A = acc[tid];
    A.assign((size_t)P * 14, 0.0);
#pragma omp for schedule(dynamic, 4)
    for (int tile = 0; tile < gx * gy; tile++) {
      const int tx = tile % gx, ty = tile / gx;
      const uint32_t rb = ranges[2 * tile], re = ranges[2 * tile + 1];
      for (int ly = 0; ly < BLOCK_Y; ly++)
        for (int lx = 0; lx < BLOCK_X; lx++) {
          const int px = tx * BLOCK_X + lx, py = ty * BLOCK_Y + ly;
          if (px >= W || py >= H) continue;
          const size_t pix_id = (size_t)W * py + px;
          const R pxf = (R)px, pyf = (R)py;
          const R T_final = final_Ts[pix_id];
          R T = T_final;
          const uint32_t last_contributor = n_contrib[pix_id];
          R accum_rec[3] = {0, 0, 0}, last_color[3] = {0, 0, 0}, dL_dpixel[3];
          R accum_rec_F[3] = {0, 0, 0}, last_F[3] = {0, 0, 0}, dL_dpixel_F[3] = {0, 0, 0};
          for (int i = 0; i < 3; i++) dL_dpixel[i] = dL_dpixels[i * N + pix_id];
          if (include_feature)
            for (int i = 0; i < 3; i++) dL_dpixel_F[i] = dL_dpixels_F[i * N + pix_id];
          R last_alpha = 0;
          uint32_t contributor = re - rb;
          for (uint32_t k = re; k-- > rb;) {
            contributor--;
            if (contributor >= last_contributor) continue;
            const uint32_t id = point_list[k];
            const R dx = means2D[2 * (size_t)id] - pxf, dy = means2D[2 * (size_t)id + 1] - pyf;
            const R* co = conic_opacity + 4 * (size_t)id;
            const R S = std::fma(dx, co[0] * dx, (co[2] * dy) * dy);
            const R power = std::fma(S, (R)-0.5, -((co[1] * dx) * dy));
            if (power > 0) continue;
            const R G = exp_(power);
            const R alpha = std::min(c_<R>(0.99), co[3] * G);
            if (alpha < (R)(1.0f / 255.0f)) continue;
            T = T / ((R)1 - alpha);
            const R dchannel_dcolor = alpha * T;
            R dL_dalpha = 0;
            double* a = A.data() + (size_t)id * 14;
            for (int ch = 0; ch < 3; ch++) {
              const R c = colors[3 * (size_t)id + ch];
              accum_rec[ch] = last_alpha * last_color[ch] + ((R)1 - last_alpha) * accum_rec[ch];
              last_color[ch] = c;
              dL_dalpha += (c - accum_rec[ch]) * dL_dpixel[ch];
              a[8 + ch] += (double)(dchannel_dcolor * dL_dpixel[ch]);
            }
            if (include_feature)
              for (int ch = 0; ch < 3; ch++) {
                const R f = lang[3 * (size_t)id + ch];
                accum_rec_F[ch] = last_alpha * last_F[ch] + ((R)1 - last_alpha) * accum_rec_F[ch];
                last_F[ch] = f;
                dL_dalpha += (f - accum_rec_F[ch]) * dL_dpixel_F[ch];
                a[11 + ch] += (double)(dchannel_dcolor * dL_dpixel_F[ch]);
              }
            dL_dalpha *= T;
            last_alpha = alpha;
            R bg_dot_dpixel = 0;
            for (int i = 0; i < 3; i++) bg_dot_dpixel += bg[i] * dL_dpixel[i];
            dL_dalpha += (-T_final / ((R)1 - alpha)) * bg_dot_dpixel;
            const R dL_dG = co[3] * dL_dalpha;
            const R gdx = G * dx, gdy = G * dy;
            const R dG_ddelx = -gdx * co[0] - gdy * co[1];
            const R dG_ddely = -gdy * co[2] - gdx * co[1];
            a[0] += (double)(dL_dG * dG_ddelx * ddelx_dx);
            a[1] += (double)(dL_dG * dG_ddely * ddely_dy);
            a[3] += (double)((R)-0.5 * gdx * dx * dL_dG);
            a[4] += (double)((R)-0.5 * gdx * dy * dL_dG);
            a[6] += (double)((R)-0.5 * gdy * dy * dL_dG);
            a[7] += (double)(G * dL_dalpha);
          }
        }
    }
  }
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; i++) {
    double s[14] = {0};
    for (int t = 0; t < nthreads; t++) {
      if (acc[t].empty()) continue;
      const double* a = acc[t].data() + (size_t)i * 14;
      for (int k = 0; k < 14; k++) s[k] += a[k];
    }
    dL_dmean2D[3 * (size_t)i + 0] += (R)s[0];
    dL_dmean2D[3 * (size_t)i + 1] += (R)s[1];
    dL_dconic[4 * (size_t)i + 0] += (R)s[3];
    dL_dconic[4 * (size_t)i + 1] += (R)s[4];
    dL_dconic[4 * (size_t)i + 3] += (R)s[6];
    dL_dopacity[i] += (R)s[7];
    for (int ch = 0; ch < 3; ch++) dL_dcolors[3 * (size_t)i + ch] += (R)s[8 + ch];
    if (include_feature)
      for (int ch = 0; ch < 3; ch++) dL_dfeat[3 * (size_t)i + ch] += (R)s[11 + ch];
  }
}

// RAST/cuda_rasterizer/auxiliary.h:107-117
template <class R>
inline void dnormvdv3(const R* v, const R* dv, R* out) {
  R sum2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
  R invsum32 = (R)1 / std::sqrt(sum2 * sum2 * sum2);
  out[0] = ((+sum2 - v[0] * v[0]) * dv[0] - v[1] * v[0] * dv[1] - v[2] * v[0] * dv[2]) * invsum32;
  out[1] = (-v[0] * v[1] * dv[0] + (sum2 - v[1] * v[1]) * dv[1] - v[2] * v[1] * dv[2]) * invsum32;
  out[2] = (-v[0] * v[2] * dv[0] - v[1] * v[2] * dv[1] + (sum2 - v[2] * v[2]) * dv[2]) * invsum32;
}

// ---------------------------------------------------------------------------------------------------
// Stage 5: per-Gaussian backward — RAST/cuda_rasterizer/backward.cu:144-274 (computeCov2DCUDA),
// :346-396 (preprocessCUDA), :20-139 (SH), :278-341 (cov3D)
// ---------------------------------------------------------------------------------------------------
template <class R>
void preprocess_bw_T(int P, int D, int M, const R* means3D, const int32_t* radii, const R* shs,
                     const uint8_t* clamped, const R* scales, const R* rotations, R scale_modifier, const R* cov3Ds,
                     const R* viewmatrix, const R* projmatrix, int W, int H, R tan_fovx, R tan_fovy, const R* campos,
                     const R* dL_dmean2D, const R* dL_dconic, R* dL_dmean3D, R* dL_dcolor, R* dL_dcov3D, R* dL_dsh,
                     R* dL_dscale, R* dL_drot) {
  const R h_y = (R)H / ((R)2 * tan_fovy);
  const R h_x = (R)W / ((R)2 * tan_fovx);
#pragma omp parallel for schedule(static)
  for (int idx = 0; idx < P; idx++) {
    if (!(radii[idx] > 0)) continue;
    const R* mean = means3D + 3 * (size_t)idx;
    // ---- computeCov2DCUDA ----
    {
      const R* cov3D = cov3Ds + 6 * (size_t)idx;
      const R dLc[3] = {dL_dconic[4 * (size_t)idx], dL_dconic[4 * (size_t)idx + 1], dL_dconic[4 * (size_t)idx + 3]};
      Cov2DCtx<R> cc;
      computeCov2D(mean, h_x, h_y, tan_fovx, tan_fovy, cov3D, viewmatrix, cc);
      const R x_grad_mul = (cc.txtz < -cc.limx || cc.txtz > cc.limx) ? 0 : 1;
      const R y_grad_mul = (cc.tytz < -cc.limy || cc.tytz > cc.limy) ? 0 : 1;
      const M3<R>&T = cc.T, &Vrk = cc.Vrk, &Wm = cc.W;
      R a = cc.cov.m[0][0], b = cc.cov.m[0][1], c = cc.cov.m[1][1];
      R denom = a * c - b * b;
      R dL_da = 0, dL_db = 0, dL_dc = 0;
      R denom2inv = (R)1 / ((denom * denom) + c_<R>(0.0000001));
      R* dcov = dL_dcov3D + 6 * (size_t)idx;
      if (denom2inv != 0) {
        dL_da = denom2inv * (-c * c * dLc[0] + 2 * b * c * dLc[1] + (denom - a * c) * dLc[2]);
        dL_dc = denom2inv * (-a * a * dLc[2] + 2 * a * b * dLc[1] + (denom - a * c) * dLc[0]);
        dL_db = denom2inv * 2 * (b * c * dLc[0] - (denom + 2 * b * b) * dLc[1] + a * b * dLc[2]);
        dcov[0] = (T.m[0][0] * T.m[0][0] * dL_da + T.m[0][0] * T.m[1][0] * dL_db + T.m[1][0] * T.m[1][0] * dL_dc);
        dcov[3] = (T.m[0][1] * T.m[0][1] * dL_da + T.m[0][1] * T.m[1][1] * dL_db + T.m[1][1] * T.m[1][1] * dL_dc);
        dcov[5] = (T.m[0][2] * T.m[0][2] * dL_da + T.m[0][2] * T.m[1][2] * dL_db + T.m[1][2] * T.m[1][2] * dL_dc);
        dcov[1] = 2 * T.m[0][0] * T.m[0][1] * dL_da + (T.m[0][0] * T.m[1][1] + T.m[0][1] * T.m[1][0]) * dL_db +
                  2 * T.m[1][0] * T.m[1][1] * dL_dc;
        dcov[2] = 2 * T.m[0][0] * T.m[0][2] * dL_da + (T.m[0][0] * T.m[1][2] + T.m[0][2] * T.m[1][0]) * dL_db +
                  2 * T.m[1][0] * T.m[1][2] * dL_dc;
        dcov[4] = 2 * T.m[0][2] * T.m[0][1] * dL_da + (T.m[0][1] * T.m[1][2] + T.m[0][2] * T.m[1][1]) * dL_db +
                  2 * T.m[1][1] * T.m[1][2] * dL_dc;
      } else {
        for (int i = 0; i < 6; i++) dcov[i] = 0;
      }
      auto row = [&](int r, int k) {  // T[r][0]*Vrk[k][0] + T[r][1]*Vrk[k][1] + T[r][2]*Vrk[k][2]
        return T.m[r][0] * Vrk.m[k][0] + T.m[r][1] * Vrk.m[k][1] + T.m[r][2] * Vrk.m[k][2];
      };
      R dL_dT00 = 2 * row(0, 0) * dL_da + row(1, 0) * dL_db;
      R dL_dT01 = 2 * row(0, 1) * dL_da + row(1, 1) * dL_db;
      R dL_dT02 = 2 * row(0, 2) * dL_da + row(1, 2) * dL_db;
      R dL_dT10 = 2 * row(1, 0) * dL_dc + row(0, 0) * dL_db;
      R dL_dT11 = 2 * row(1, 1) * dL_dc + row(0, 1) * dL_db;
      R dL_dT12 = 2 * row(1, 2) * dL_dc + row(0, 2) * dL_db;
      R dL_dJ00 = Wm.m[0][0] * dL_dT00 + Wm.m[0][1] * dL_dT01 + Wm.m[0][2] * dL_dT02;
      R dL_dJ02 = Wm.m[2][0] * dL_dT00 + Wm.m[2][1] * dL_dT01 + Wm.m[2][2] * dL_dT02;
      R dL_dJ11 = Wm.m[1][0] * dL_dT10 + Wm.m[1][1] * dL_dT11 + Wm.m[1][2] * dL_dT12;
      R dL_dJ12 = Wm.m[2][0] * dL_dT10 + Wm.m[2][1] * dL_dT11 + Wm.m[2][2] * dL_dT12;
      R tz = (R)1 / cc.tz, tz2 = tz * tz, tz3 = tz2 * tz;
      R dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
      R dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
      R dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * cc.tx) * tz3 * dL_dJ02 +
                 (2 * h_y * cc.ty) * tz3 * dL_dJ12;
      const R* V = viewmatrix;  // transformVec4x3Transpose
      dL_dmean3D[3 * (size_t)idx + 0] = V[0] * dL_dtx + V[1] * dL_dty + V[2] * dL_dtz;
      dL_dmean3D[3 * (size_t)idx + 1] = V[4] * dL_dtx + V[5] * dL_dty + V[6] * dL_dtz;
      dL_dmean3D[3 * (size_t)idx + 2] = V[8] * dL_dtx + V[9] * dL_dty + V[10] * dL_dtz;
    }
    // ---- preprocessCUDA (backward) ----
    {
      const R* proj = projmatrix;
      R m[3] = {mean[0], mean[1], mean[2]};
      R m_hw = proj[3] * m[0] + proj[7] * m[1] + proj[11] * m[2] + proj[15];
      R m_w = (R)1 / (m_hw + c_<R>(0.0000001));
      R mul1 = (proj[0] * m[0] + proj[4] * m[1] + proj[8] * m[2] + proj[12]) * m_w * m_w;
      R mul2 = (proj[1] * m[0] + proj[5] * m[1] + proj[9] * m[2] + proj[13]) * m_w * m_w;
      const R gx_ = dL_dmean2D[3 * (size_t)idx], gy_ = dL_dmean2D[3 * (size_t)idx + 1];
      dL_dmean3D[3 * (size_t)idx + 0] += (proj[0] * m_w - proj[3] * mul1) * gx_ + (proj[1] * m_w - proj[3] * mul2) * gy_;
      dL_dmean3D[3 * (size_t)idx + 1] += (proj[4] * m_w - proj[7] * mul1) * gx_ + (proj[5] * m_w - proj[7] * mul2) * gy_;
      dL_dmean3D[3 * (size_t)idx + 2] += (proj[8] * m_w - proj[11] * mul1) * gx_ + (proj[9] * m_w - proj[11] * mul2) * gy_;
    }
    if (shs) {  // computeColorFromSH backward, backward.cu:20-139
      const R* sh = shs + (size_t)idx * M * 3;
      R* dsh = dL_dsh + (size_t)idx * M * 3;
      R dir_orig[3] = {mean[0] - campos[0], mean[1] - campos[1], mean[2] - campos[2]};
      R len = std::sqrt(dir_orig[0] * dir_orig[0] + dir_orig[1] * dir_orig[1] + dir_orig[2] * dir_orig[2]);
      R x = dir_orig[0] / len, y = dir_orig[1] / len, z = dir_orig[2] / len;
      R dL_dRGB[3];
      for (int ch = 0; ch < 3; ch++) dL_dRGB[ch] = dL_dcolor[3 * (size_t)idx + ch] * (clamped[3 * (size_t)idx + ch] ? 0 : 1);
      R dRGBdx[3] = {0, 0, 0}, dRGBdy[3] = {0, 0, 0}, dRGBdz[3] = {0, 0, 0};
      auto S = [&](int k, int ch) { return sh[k * 3 + ch]; };
      auto setsh = [&](int k, R w) { for (int ch = 0; ch < 3; ch++) dsh[k * 3 + ch] = w * dL_dRGB[ch]; };
      setsh(0, c_<R>(SH_C0));
      if (D > 0) {
        setsh(1, -c_<R>(SH_C1) * y);
        setsh(2, c_<R>(SH_C1) * z);
        setsh(3, -c_<R>(SH_C1) * x);
        for (int ch = 0; ch < 3; ch++) {
          dRGBdx[ch] = -c_<R>(SH_C1) * S(3, ch);
          dRGBdy[ch] = -c_<R>(SH_C1) * S(1, ch);
          dRGBdz[ch] = c_<R>(SH_C1) * S(2, ch);
        }
        if (D > 1) {
          R xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
          setsh(4, c_<R>(SH_C2[0]) * xy);
          setsh(5, c_<R>(SH_C2[1]) * yz);
          setsh(6, c_<R>(SH_C2[2]) * ((R)2 * zz - xx - yy));
          setsh(7, c_<R>(SH_C2[3]) * xz);
          setsh(8, c_<R>(SH_C2[4]) * (xx - yy));
          for (int ch = 0; ch < 3; ch++) {
            dRGBdx[ch] += c_<R>(SH_C2[0]) * y * S(4, ch) + c_<R>(SH_C2[2]) * (R)2 * -x * S(6, ch) +
                          c_<R>(SH_C2[3]) * z * S(7, ch) + c_<R>(SH_C2[4]) * (R)2 * x * S(8, ch);
            dRGBdy[ch] += c_<R>(SH_C2[0]) * x * S(4, ch) + c_<R>(SH_C2[1]) * z * S(5, ch) +
                          c_<R>(SH_C2[2]) * (R)2 * -y * S(6, ch) + c_<R>(SH_C2[4]) * (R)2 * -y * S(8, ch);
            dRGBdz[ch] += c_<R>(SH_C2[1]) * y * S(5, ch) + c_<R>(SH_C2[2]) * (R)2 * (R)2 * z * S(6, ch) +
                          c_<R>(SH_C2[3]) * x * S(7, ch);
          }
          if (D > 2) {
            setsh(9, c_<R>(SH_C3[0]) * y * ((R)3 * xx - yy));
            setsh(10, c_<R>(SH_C3[1]) * xy * z);
            setsh(11, c_<R>(SH_C3[2]) * y * ((R)4 * zz - xx - yy));
            setsh(12, c_<R>(SH_C3[3]) * z * ((R)2 * zz - (R)3 * xx - (R)3 * yy));
            setsh(13, c_<R>(SH_C3[4]) * x * ((R)4 * zz - xx - yy));
            setsh(14, c_<R>(SH_C3[5]) * z * (xx - yy));
            setsh(15, c_<R>(SH_C3[6]) * x * (xx - (R)3 * yy));
            for (int ch = 0; ch < 3; ch++) {
              dRGBdx[ch] += (c_<R>(SH_C3[0]) * S(9, ch) * (R)3 * (R)2 * xy + c_<R>(SH_C3[1]) * S(10, ch) * yz +
                             c_<R>(SH_C3[2]) * S(11, ch) * (R)-2 * xy + c_<R>(SH_C3[3]) * S(12, ch) * (R)-3 * (R)2 * xz +
                             c_<R>(SH_C3[4]) * S(13, ch) * ((R)-3 * xx + (R)4 * zz - yy) +
                             c_<R>(SH_C3[5]) * S(14, ch) * (R)2 * xz + c_<R>(SH_C3[6]) * S(15, ch) * (R)3 * (xx - yy));
              dRGBdy[ch] += (c_<R>(SH_C3[0]) * S(9, ch) * (R)3 * (xx - yy) + c_<R>(SH_C3[1]) * S(10, ch) * xz +
                             c_<R>(SH_C3[2]) * S(11, ch) * ((R)-3 * yy + (R)4 * zz - xx) +
                             c_<R>(SH_C3[3]) * S(12, ch) * (R)-3 * (R)2 * yz + c_<R>(SH_C3[4]) * S(13, ch) * (R)-2 * xy +
                             c_<R>(SH_C3[5]) * S(14, ch) * (R)-2 * yz + c_<R>(SH_C3[6]) * S(15, ch) * (R)-3 * (R)2 * xy);
              dRGBdz[ch] += (c_<R>(SH_C3[1]) * S(10, ch) * xy + c_<R>(SH_C3[2]) * S(11, ch) * (R)4 * (R)2 * yz +
                             c_<R>(SH_C3[3]) * S(12, ch) * (R)3 * ((R)2 * zz - xx - yy) +
                             c_<R>(SH_C3[4]) * S(13, ch) * (R)4 * (R)2 * xz + c_<R>(SH_C3[5]) * S(14, ch) * (xx - yy));
            }
          }
        }
      }
      R dL_ddir[3] = {0, 0, 0};
      for (int ch = 0; ch < 3; ch++) {
        dL_ddir[0] += dRGBdx[ch] * dL_dRGB[ch];
        dL_ddir[1] += dRGBdy[ch] * dL_dRGB[ch];
        dL_ddir[2] += dRGBdz[ch] * dL_dRGB[ch];
      }
      R dm[3];
      dnormvdv3(dir_orig, dL_ddir, dm);
      for (int k = 0; k < 3; k++) dL_dmean3D[3 * (size_t)idx + k] += dm[k];
    }
    if (scales) {  // computeCov3D backward, backward.cu:278-341
      const R* rot = rotations + 4 * (size_t)idx;
      const R* scale = scales + 3 * (size_t)idx;
      R r = rot[0], x = rot[1], y = rot[2], z = rot[3];
      M3<R> Rm = quatR(rot);
      R s[3] = {scale_modifier * scale[0], scale_modifier * scale[1], scale_modifier * scale[2]};
      M3<R> Mm;
      for (int c = 0; c < 3; c++)
        for (int rr = 0; rr < 3; rr++) Mm.m[c][rr] = s[rr] * Rm.m[c][rr];
      const R* g = dL_dcov3D + 6 * (size_t)idx;
      M3<R> dSig;
      dSig.m[0][0] = g[0]; dSig.m[0][1] = (R)0.5 * g[1]; dSig.m[0][2] = (R)0.5 * g[2];
      dSig.m[1][0] = (R)0.5 * g[1]; dSig.m[1][1] = g[3]; dSig.m[1][2] = (R)0.5 * g[4];
      dSig.m[2][0] = (R)0.5 * g[2]; dSig.m[2][1] = (R)0.5 * g[4]; dSig.m[2][2] = g[5];
      M3<R> M2;
      for (int c = 0; c < 3; c++)
        for (int rr = 0; rr < 3; rr++) M2.m[c][rr] = (R)2 * Mm.m[c][rr];
      M3<R> dL_dM = mul(M2, dSig);
      M3<R> Rt = transpose(Rm), dMt = transpose(dL_dM);
      R* dsc = dL_dscale + 3 * (size_t)idx;
      for (int k = 0; k < 3; k++)
        dsc[k] = Rt.m[k][0] * dMt.m[k][0] + Rt.m[k][1] * dMt.m[k][1] + Rt.m[k][2] * dMt.m[k][2];
      for (int k = 0; k < 3; k++)
        for (int rr = 0; rr < 3; rr++) dMt.m[k][rr] *= s[k];
      R* dq = dL_drot + 4 * (size_t)idx;
      dq[0] = 2 * z * (dMt.m[0][1] - dMt.m[1][0]) + 2 * y * (dMt.m[2][0] - dMt.m[0][2]) + 2 * x * (dMt.m[1][2] - dMt.m[2][1]);
      dq[1] = 2 * y * (dMt.m[1][0] + dMt.m[0][1]) + 2 * z * (dMt.m[2][0] + dMt.m[0][2]) + 2 * r * (dMt.m[1][2] - dMt.m[2][1]) -
              4 * x * (dMt.m[2][2] + dMt.m[1][1]);
      dq[2] = 2 * x * (dMt.m[1][0] + dMt.m[0][1]) + 2 * r * (dMt.m[2][0] - dMt.m[0][2]) + 2 * z * (dMt.m[1][2] + dMt.m[2][1]) -
              4 * y * (dMt.m[2][2] + dMt.m[0][0]);
      dq[3] = 2 * r * (dMt.m[0][1] - dMt.m[1][0]) + 2 * x * (dMt.m[2][0] + dMt.m[0][2]) + 2 * y * (dMt.m[1][2] + dMt.m[2][1]) -
              4 * z * (dMt.m[1][1] + dMt.m[0][0]);
    }
  }
}

// RAST/cuda_rasterizer/rasterizer_impl.cu:54-66
template <class R>
void mark_visible_T(int P, const R* means3D, const R* viewmatrix, uint8_t* present) {
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; i++) {
    const R* p = means3D + 3 * (size_t)i;
    present[i] = tp_row(viewmatrix, 2, p[0], p[1], p[2]) > c_<R>(0.2);
  }
}

}  // namespace

#define ORC_EXPORT extern "C" __attribute__((visibility("default")))

#define ORC_DEFINE(SUF, R)                                                                                              \
  ORC_EXPORT int64_t orc_preprocess_##SUF(                                                                              \
      int P, int D, int M, const R* means3D, const R* scales, R scale_modifier, const R* rotations, const R* opacities, \
      const R* shs, const R* cov3D_precomp, const R* colors_precomp, const R* viewmatrix, const R* projmatrix,          \
      const R* campos, int W, int H, R tan_fovx, R tan_fovy, int32_t* radii, R* means2D, R* depths, R* cov3Ds, R* rgb,  \
      R* conic_opacity, uint8_t* clamped, uint32_t* tiles_touched, uint32_t* point_offsets) {                           \
    return preprocess_T<R>(P, D, M, means3D, scales, scale_modifier, rotations, opacities, shs, cov3D_precomp,          \
                           colors_precomp, viewmatrix, projmatrix, campos, W, H, tan_fovx, tan_fovy, radii, means2D,    \
                           depths, cov3Ds, rgb, conic_opacity, clamped, tiles_touched, point_offsets);                  \
  }                                                                                                                     \
  ORC_EXPORT void orc_binning_##SUF(int P, int W, int H, const int32_t* radii, const R* means2D, const R* depths,       \
                                    const uint32_t* point_offsets, int64_t num_rendered, uint64_t* keys_unsorted,       \
                                    uint32_t* values_unsorted, uint64_t* keys_sorted, uint32_t* point_list,             \
                                    uint32_t* ranges) {                                                                 \
    binning_T<R>(P, W, H, radii, means2D, depths, point_offsets, num_rendered, keys_unsorted, values_unsorted,          \
                 keys_sorted, point_list, ranges);                                                                      \
  }                                                                                                                     \
  ORC_EXPORT void orc_render_##SUF(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const R* means2D,  \
                                   const R* colors, const R* lang, const R* conic_opacity, const R* bg,                 \
                                   int include_feature, R* final_T, uint32_t* n_contrib, R* out_color, R* out_feature,  \
                                   int64_t* k_exam, int64_t* k_contrib) {                                               \
    render_T<R>(W, H, ranges, point_list, means2D, colors, lang, conic_opacity, bg, include_feature, final_T,           \
                n_contrib, out_color, out_feature, k_exam, k_contrib);                                                  \
  }                                                                                                                     \
  ORC_EXPORT void orc_render_bw_##SUF(int P, int W, int H, const uint32_t* ranges, const uint32_t* point_list,          \
                                      const R* bg, const R* means2D, const R* conic_opacity, const R* colors,           \
                                      const R* lang, const R* final_Ts, const uint32_t* n_contrib,                      \
                                      const R* dL_dpixels, const R* dL_dpixels_F, int include_feature, R* dL_dmean2D,   \
                                      R* dL_dconic, R* dL_dopacity, R* dL_dcolors, R* dL_dfeat) {                       \
    render_bw_T<R>(P, W, H, ranges, point_list, bg, means2D, conic_opacity, colors, lang, final_Ts, n_contrib,          \
                   dL_dpixels, dL_dpixels_F, include_feature, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolors,           \
                   dL_dfeat);                                                                                           \
  }                                                                                                                     \
  ORC_EXPORT void orc_preprocess_bw_##SUF(                                                                              \
      int P, int D, int M, const R* means3D, const int32_t* radii, const R* shs, const uint8_t* clamped,                \
      const R* scales, const R* rotations, R scale_modifier, const R* cov3Ds, const R* viewmatrix,                      \
      const R* projmatrix, int W, int H, R tan_fovx, R tan_fovy, const R* campos, const R* dL_dmean2D,                  \
      const R* dL_dconic, R* dL_dmean3D, R* dL_dcolor, R* dL_dcov3D, R* dL_dsh, R* dL_dscale, R* dL_drot) {             \
    preprocess_bw_T<R>(P, D, M, means3D, radii, shs, clamped, scales, rotations, scale_modifier, cov3Ds, viewmatrix,    \
                       projmatrix, W, H, tan_fovx, tan_fovy, campos, dL_dmean2D, dL_dconic, dL_dmean3D, dL_dcolor,      \
                       dL_dcov3D, dL_dsh, dL_dscale, dL_drot);                                                          \
  }                                                                                                                     \
  ORC_EXPORT void orc_mark_visible_##SUF(int P, const R* means3D, const R* viewmatrix, uint8_t* present) {              \
    mark_visible_T<R>(P, means3D, viewmatrix, present);                                                                 \
  }

ORC_DEFINE(f32, float)
ORC_DEFINE(f64, double)

ORC_EXPORT uint32_t orc_higher_msb(uint32_t n) { return getHigherMsb(n); }
ORC_EXPORT int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
ORC_EXPORT void orc_set_num_threads(int n) {
#ifdef _OPENMP
  omp_set_num_threads(n);
#else
  (void)n;
#endif
}
