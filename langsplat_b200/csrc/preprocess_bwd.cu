// preprocess_bwd.cu — K8'+K9' fused: per-Gaussian backward through conic -> cov2D -> cov3D -> (scale, rotation),
// through the projection (mean2D -> mean3D) and through the SH colour.
// Replaces computeCov2DCUDA (RAST/cuda_rasterizer/backward.cu:144-274), preprocessCUDA (backward.cu:346-396),
// computeColorFromSH backward (:20-139) and computeCov3D backward (:278-341), which the reference runs as two
// kernels with dL_dcov3D / dL_dmean3D round-tripping through global memory, after zero-filling ten tensors
// (RAST/rasterize_points.cu:166-183).  Here one kernel writes every output element exactly once (zeros for
// Gaussians with radius 0), so the caller can hand in uninitialised memory.  HBM-bound.
// With `moments` set, the blend kernel (render_bwd.cu) left the pair-weighted moments M0, (Mx, My), (Mxx, Mxy, Myy) of the
// pixel offsets in dL_dopacity / dL_dmean2D / dL_dconic instead of the reference's sums (backward.cu:552-582); the
// per-Gaussian coefficients are applied here, and dL_dmean2D is rewritten in place with the actual gradient.
#include <algorithm>
#include <vector>

#include "common.cuh"
#include "gs_math.cuh"
#include "stage.cuh"

namespace lsr {

namespace {

__device__ __forceinline__ float3 dnormvdv(float3 v, float3 dv) {  // RAST/cuda_rasterizer/auxiliary.h:107-117
  const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
  const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
  float3 o;
  o.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
  o.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
  o.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
  return o;
}

// SH backward: writes dL_dsh rows [0,(D+1)^2), zero for the rest up to M, returns dL/d(mean) through the direction.
template <bool WRITE>
__device__ __forceinline__ float3 sh_backward_t(int deg, int M, const float* __restrict__ sh, float3 dir_orig,
                                                float3 dL_dRGB, float* __restrict__ dsh) {
  const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
  const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
  const float g[3] = {dL_dRGB.x, dL_dRGB.y, dL_dRGB.z};
  float dRGBdx[3] = {0, 0, 0}, dRGBdy[3] = {0, 0, 0}, dRGBdz[3] = {0, 0, 0};
  auto put = [&](int k, float w) {
    if (!WRITE) return;
    dsh[k * 3 + 0] = w * g[0];
    dsh[k * 3 + 1] = w * g[1];
    dsh[k * 3 + 2] = w * g[2];
  };
  int written = 1;
  put(0, LSR_SH_C0);
  if (deg > 0) {
    put(1, -LSR_SH_C1 * y);
    put(2, LSR_SH_C1 * z);
    put(3, -LSR_SH_C1 * x);
    written = 4;
#pragma unroll
    for (int ch = 0; ch < 3; ch++) {
      dRGBdx[ch] = -LSR_SH_C1 * sh[3 * 3 + ch];
      dRGBdy[ch] = -LSR_SH_C1 * sh[1 * 3 + ch];
      dRGBdz[ch] = LSR_SH_C1 * sh[2 * 3 + ch];
    }
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      put(4, LSR_SH_C2_0 * xy);
      put(5, LSR_SH_C2_1 * yz);
      put(6, LSR_SH_C2_2 * (2.f * zz - xx - yy));
      put(7, LSR_SH_C2_3 * xz);
      put(8, LSR_SH_C2_4 * (xx - yy));
      written = 9;
#pragma unroll
      for (int ch = 0; ch < 3; ch++) {
        const float s4 = sh[4 * 3 + ch], s5 = sh[5 * 3 + ch], s6 = sh[6 * 3 + ch], s7 = sh[7 * 3 + ch],
                    s8 = sh[8 * 3 + ch];
        dRGBdx[ch] += LSR_SH_C2_0 * y * s4 + LSR_SH_C2_2 * 2.f * -x * s6 + LSR_SH_C2_3 * z * s7 +
                      LSR_SH_C2_4 * 2.f * x * s8;
        dRGBdy[ch] += LSR_SH_C2_0 * x * s4 + LSR_SH_C2_1 * z * s5 + LSR_SH_C2_2 * 2.f * -y * s6 +
                      LSR_SH_C2_4 * 2.f * -y * s8;
        dRGBdz[ch] += LSR_SH_C2_1 * y * s5 + LSR_SH_C2_2 * 2.f * 2.f * z * s6 + LSR_SH_C2_3 * x * s7;
      }
      if (deg > 2) {
        put(9, LSR_SH_C3_0 * y * (3.f * xx - yy));
        put(10, LSR_SH_C3_1 * xy * z);
        put(11, LSR_SH_C3_2 * y * (4.f * zz - xx - yy));
        put(12, LSR_SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy));
        put(13, LSR_SH_C3_4 * x * (4.f * zz - xx - yy));
        put(14, LSR_SH_C3_5 * z * (xx - yy));
        put(15, LSR_SH_C3_6 * x * (xx - 3.f * yy));
        written = 16;
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
          const float s9 = sh[9 * 3 + ch], s10 = sh[10 * 3 + ch], s11 = sh[11 * 3 + ch], s12 = sh[12 * 3 + ch],
                      s13 = sh[13 * 3 + ch], s14 = sh[14 * 3 + ch], s15 = sh[15 * 3 + ch];
          dRGBdx[ch] += (LSR_SH_C3_0 * s9 * 3.f * 2.f * xy + LSR_SH_C3_1 * s10 * yz + LSR_SH_C3_2 * s11 * -2.f * xy +
                         LSR_SH_C3_3 * s12 * -3.f * 2.f * xz + LSR_SH_C3_4 * s13 * (-3.f * xx + 4.f * zz - yy) +
                         LSR_SH_C3_5 * s14 * 2.f * xz + LSR_SH_C3_6 * s15 * 3.f * (xx - yy));
          dRGBdy[ch] += (LSR_SH_C3_0 * s9 * 3.f * (xx - yy) + LSR_SH_C3_1 * s10 * xz +
                         LSR_SH_C3_2 * s11 * (-3.f * yy + 4.f * zz - xx) + LSR_SH_C3_3 * s12 * -3.f * 2.f * yz +
                         LSR_SH_C3_4 * s13 * -2.f * xy + LSR_SH_C3_5 * s14 * -2.f * yz +
                         LSR_SH_C3_6 * s15 * -3.f * 2.f * xy);
          dRGBdz[ch] += (LSR_SH_C3_1 * s10 * xy + LSR_SH_C3_2 * s11 * 4.f * 2.f * yz +
                         LSR_SH_C3_3 * s12 * 3.f * (2.f * zz - xx - yy) + LSR_SH_C3_4 * s13 * 4.f * 2.f * xz +
                         LSR_SH_C3_5 * s14 * (xx - yy));
        }
      }
    }
  }
  if (WRITE)
    for (int k = written; k < M; k++) put(k, 0.f);  // entries above the active degree stay zero
  float3 dL_ddir;
  dL_ddir.x = dRGBdx[0] * g[0] + dRGBdx[1] * g[1] + dRGBdx[2] * g[2];
  dL_ddir.y = dRGBdy[0] * g[0] + dRGBdy[1] * g[1] + dRGBdy[2] * g[2];
  dL_ddir.z = dRGBdz[0] * g[0] + dRGBdz[1] * g[1] + dRGBdz[2] * g[2];
  return dnormvdv(dir_orig, dL_ddir);
}

__device__ __forceinline__ float3 sh_backward(int deg, int M, const float* __restrict__ sh, float3 dir_orig,
                                              float3 dL_dRGB, float* __restrict__ dsh) {
  return sh_backward_t<true>(deg, M, sh, dir_orig, dL_dRGB, dsh);
}
// direction gradient only (the dL_dsh rows are produced from sh_basis by the caller)
__device__ __forceinline__ float3 sh_dir_backward(int deg, const float* sh, float3 dir_orig, const float (&g)[3]) {
  return sh_backward_t<false>(deg, 16, sh, dir_orig, make_float3(g[0], g[1], g[2]), nullptr);
}

constexpr int BPRE_THREADS = 256;
constexpr int SH_ROW = 12;
constexpr int BPRE_WARP_F4 = 32 * (SH_ROW + 1) + 48;  // per warp: 32 padded SH rows + 32 cov3D rows (24 B each)
constexpr size_t BPRE_SH_SMEM = (size_t)(BPRE_THREADS / 32) * BPRE_WARP_F4 * sizeof(float4);

// SH16: shs / dL_dsh are [P,16,3], 16-byte aligned -> rows move through shared memory as coalesced 16-byte chunks.
template <bool SH16>
__global__ void __launch_bounds__(BPRE_THREADS, 3)
preprocess_bwd_kernel(int P, int D, int M, const float* __restrict__ means3D, const int32_t* __restrict__ radii,
                      const float* __restrict__ shs, const uint8_t* __restrict__ clamped,
                      const float* __restrict__ scales, const float* __restrict__ rotations, float scale_modifier,
                      const float* __restrict__ cov3Ds, const float* __restrict__ viewmatrix,
                      const float* __restrict__ projmatrix, float h_x, float h_y, float tan_fovx, float tan_fovy,
                      const float* __restrict__ campos, const float* __restrict__ opacities, int moments, float ddelx_dx,
                      float ddely_dy, float* __restrict__ dL_dmean2D,
                      const float* __restrict__ dL_dconic, const float* __restrict__ dL_dcolor,
                      float* __restrict__ dL_dmean3D, float* __restrict__ dL_dcov3D, float* __restrict__ dL_dsh,
                      float* __restrict__ dL_dscale, float* __restrict__ dL_drot, float* __restrict__ dL_dcolor_masked,
                      const float4* __restrict__ acc, float* __restrict__ dL_dopacity_out,
                      float* __restrict__ dL_dcolor_out, float* __restrict__ dL_dfeat_out) {
  // `acc` (lsr_grads::blend_acc): the blend kernel left everything it accumulated in one 64-byte record per Gaussian
  // instead of dL_dmean2D / dL_dconic / dL_dcolor / dL_dopacity / dL_dlanguage_feature; this kernel then reads the
  // record and WRITES those tensors (every element: zeros for Gaussians nobody touched)
  extern __shared__ __align__(16) float4 s_dyn[];
  __shared__ __align__(16) float s_xyz[BPRE_THREADS * 3];   // means3D in, dL_dmean3D out
  __shared__ __align__(16) float s_scl[BPRE_THREADS * 3];   // scales in, dL_dscale out
  __shared__ __align__(16) float s_m2d[BPRE_THREADS * 3];   // dL_dmean2D in
  __shared__ __align__(16) float s_dcol[BPRE_THREADS * 3];  // dL_dcolor in
  __shared__ float s_view[16], s_proj[16], s_cam[3];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int first = blockIdx.x * BPRE_THREADS;
  const int count = min(BPRE_THREADS, P - first);
  if (tid < 16) {
    s_view[tid] = viewmatrix[tid];
    s_proj[tid] = projmatrix[tid];
  }
  if (tid < 3) s_cam[tid] = campos[tid];
  // factored SH gradient: dL_dsh[k][ch] = basis_k(dir) * g[ch] with g = dL_dcolor masked by the clamp (backward.cu:31-34)
  // — a caller that wants the factor g instead of the 16x larger product passes dL_dcolor_masked and no dL_dsh
  const bool write_sh = dL_dsh != nullptr;
  const bool do_sh = shs != nullptr && (write_sh || dL_dcolor_masked != nullptr) && (dL_dcolor != nullptr || acc != nullptr);
  const bool do_scale = scales != nullptr && dL_dscale != nullptr;
  const bool valid = tid < count;
  const int idx = first + tid;
  const bool vis = valid && radii[idx] > 0;
  const uint32_t vis_mask = __ballot_sync(0xffffffffu, vis);
  const uint32_t valid_mask = __ballot_sync(0xffffffffu, valid);
  float4* wsm = s_dyn + (size_t)warp * BPRE_WARP_F4;
  float2* wcov = reinterpret_cast<float2*>(wsm + 32 * (SH_ROW + 1));
  const size_t wfirst = (size_t)(first + warp * 32);

  // Everything this CTA reads is requested up front with cp.async: group 0 = the flat triplet arrays and the cov3D
  // rows (needed first), group 1 = the SH rows of the visible Gaussians, which then fly during the geometry backward.
  stage_flat_async<BPRE_THREADS>(means3D, s_xyz, (size_t)first * 3, count * 3, tid);
  if (!acc) stage_flat_async<BPRE_THREADS>(dL_dmean2D, s_m2d, (size_t)first * 3, count * 3, tid);
  if (do_scale) stage_flat_async<BPRE_THREADS>(scales, s_scl, (size_t)first * 3, count * 3, tid);
  if (do_sh && !acc) stage_flat_async<BPRE_THREADS>(dL_dcolor, s_dcol, (size_t)first * 3, count * 3, tid);
  float4 rec_m = make_float4(0.f, 0.f, 0.f, 0.f), rec_q = rec_m, rec_c = rec_m, rec_f = rec_m;
  if (acc && valid) {  // this thread's record: four 16-byte loads, in flight beside the cp.async groups
    rec_m = __ldg(acc + 4 * (size_t)idx);
    rec_q = __ldg(acc + 4 * (size_t)idx + 1);
    rec_c = __ldg(acc + 4 * (size_t)idx + 2);
    rec_f = __ldg(acc + 4 * (size_t)idx + 3);
  }
  {
    const float2* src = reinterpret_cast<const float2*>(cov3Ds + wfirst * 6);
#pragma unroll
    for (int it = 0; it < 3; it++) {
      const int c = it * 32 + lane;
      if ((vis_mask >> (c / 3)) & 1u) cp_async8(wcov + c, src + c);
    }
  }
  cp_async_commit();
  if (do_sh && SH16)
    warp_load_rows_async<SH_ROW>(reinterpret_cast<const float4*>(shs) + wfirst * SH_ROW, wsm, vis_mask, lane);
  cp_async_commit();
  cp_async_wait<1>();
  __syncthreads();
  if (acc) {
    s_m2d[3 * tid] = rec_m.x; s_m2d[3 * tid + 1] = rec_m.y; s_m2d[3 * tid + 2] = 0.f;
    s_dcol[3 * tid] = rec_c.x; s_dcol[3 * tid + 1] = rec_c.y; s_dcol[3 * tid + 2] = rec_c.z;
    if (valid) {
      if (dL_dopacity_out) dL_dopacity_out[idx] = rec_m.z;  // M0 = sum of G * dL_dalpha IS dL_dopacity
      if (dL_dfeat_out) {
        dL_dfeat_out[3 * (size_t)idx] = rec_f.x; dL_dfeat_out[3 * (size_t)idx + 1] = rec_f.y;
        dL_dfeat_out[3 * (size_t)idx + 2] = rec_f.z;
      }
      if (dL_dcolor_out) {
        dL_dcolor_out[3 * (size_t)idx] = rec_c.x; dL_dcolor_out[3 * (size_t)idx + 1] = rec_c.y;
        dL_dcolor_out[3 * (size_t)idx + 2] = rec_c.z;
      }
    }
  }
  float cov3D[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (vis) {
    const float2 a = wcov[lane * 3], b = wcov[lane * 3 + 1], c = wcov[lane * 3 + 2];
    cov3D[0] = a.x; cov3D[1] = a.y; cov3D[2] = b.x; cov3D[3] = b.y; cov3D[4] = c.x; cov3D[5] = c.y;
  }
  __syncwarp();
  const float mx = s_xyz[3 * tid], my = s_xyz[3 * tid + 1], mz = s_xyz[3 * tid + 2];
  float dcov[6] = {0, 0, 0, 0, 0, 0};
  float3 dmean = make_float3(0.f, 0.f, 0.f);
  if (vis) {
    // ---------------- computeCov2DCUDA (backward.cu:144-274) ----------------
    Cov2D c2;
    compute_cov2D(mx, my, mz, h_x, h_y, tan_fovx, tan_fovy, cov3D, s_view, c2);
    const float4 dcon = acc ? make_float4(rec_q.x, rec_q.y, 0.f, rec_q.z)
                            : __ldg(reinterpret_cast<const float4*>(dL_dconic) + idx);
    float dLc_x = dcon.x, dLc_y = dcon.y, dLc_z = dcon.w;
    if (moments) {
      // dL_dconic = -0.5 * o * (Mxx, Mxy, Myy);  dL_dmean2D = -o * (cx*Mx + cy*My, cz*My + cy*Mx) * (W/2, H/2).
      // The conic is re-derived from the covariance just computed, exactly as the forward does (forward.cu:219-224).
      const float det = __fmaf_rn(c2.a, c2.c, -__fmul_rn(c2.b, c2.b));
      const float det_inv = __fdiv_rn(1.f, det);
      const float cx = __fmul_rn(c2.c, det_inv), cy = __fmul_rn(-c2.b, det_inv), cz = __fmul_rn(c2.a, det_inv);
      const float o = __ldg(opacities + idx);
      const float mo = -0.5f * o;
      dLc_x *= mo; dLc_y *= mo; dLc_z *= mo;
      const float Mx = s_m2d[3 * tid], My = s_m2d[3 * tid + 1];
      s_m2d[3 * tid] = -o * (cx * Mx + cy * My) * ddelx_dx;
      s_m2d[3 * tid + 1] = -o * (cz * My + cy * Mx) * ddely_dy;
    }
    // Reverse mode through  conic = Sigma^-1,  Sigma = (Tm Vrk Tm^T) + 0.3 I,  Tm = the two rows of the projected
    // Jacobian (W J)^T that matter,  J = J(t),  t = view * mean  — written as the matrix identities they are
    // (the reference spells the same chain out entry by entry, backward.cu:196-270):
    //   inverse:    dL/dSigma = -adj(Sigma) G adj(Sigma) / (det^2 + 1e-7),  G = [[g_xx, g_xy], [g_xy, g_yy]] the
    //               gradient w.r.t. the conic's three stored entries (the regulariser is the reference's);
    //   congruence: dL/dVrk = Tm^T D Tm,  dL/dTm = 2 D Tm Vrk,  D = dL/dSigma as a symmetric matrix;
    //   product:    dL/dJ = dL/dTm W^T  (only J's four non-constant entries are read off).
    const Mat3& Tm = c2.T;
    const float sa = c2.a, sb = c2.b, sc = c2.c;
    const float det = sa * sc - sb * sb;
    const float scale = 1.0f / (det * det + 0.0000001f);
    float D[2][2] = {{0.f, 0.f}, {0.f, 0.f}};  // dL/dSigma
    if (scale != 0) {
      const float adj[2][2] = {{sc, -sb}, {-sb, sa}};
      const float G[2][2] = {{dLc_x, dLc_y}, {dLc_y, dLc_z}};
      float GA[2][2];
#pragma unroll
      for (int r = 0; r < 2; r++)
#pragma unroll
        for (int q = 0; q < 2; q++) GA[r][q] = G[r][0] * adj[0][q] + G[r][1] * adj[1][q];
#pragma unroll
      for (int r = 0; r < 2; r++)
#pragma unroll
        for (int q = r; q < 2; q++) D[r][q] = D[q][r] = -scale * (adj[r][0] * GA[0][q] + adj[r][1] * GA[1][q]);
      // dL/dVrk = Tm^T D Tm; the six stored entries of a symmetric matrix, off-diagonal ones counted twice
      float DT[2][3];
#pragma unroll
      for (int r = 0; r < 2; r++)
#pragma unroll
        for (int k = 0; k < 3; k++) DT[r][k] = D[r][0] * Tm.m[0][k] + D[r][1] * Tm.m[1][k];
      auto vrk_grad = [&](int i, int j) { return Tm.m[0][i] * DT[0][j] + Tm.m[1][i] * DT[1][j]; };
      dcov[0] = vrk_grad(0, 0);
      dcov[1] = 2.f * vrk_grad(0, 1);
      dcov[2] = 2.f * vrk_grad(0, 2);
      dcov[3] = vrk_grad(1, 1);
      dcov[4] = 2.f * vrk_grad(1, 2);
      dcov[5] = vrk_grad(2, 2);
    }
    // dL/dTm = 2 D (Tm Vrk)
    float TV[2][3], dTm[2][3];
#pragma unroll
    for (int r = 0; r < 2; r++)
#pragma unroll
      for (int k = 0; k < 3; k++)
        TV[r][k] = Tm.m[r][0] * c2.Vrk.m[0][k] + Tm.m[r][1] * c2.Vrk.m[1][k] + Tm.m[r][2] * c2.Vrk.m[2][k];
#pragma unroll
    for (int r = 0; r < 2; r++)
#pragma unroll
      for (int k = 0; k < 3; k++) dTm[r][k] = 2.f * (D[r][0] * TV[0][k] + D[r][1] * TV[1][k]);
    // dL/dJ = dL/dTm W^T, read off at J = [[fx/z, 0, -fx x/z^2], [0, fy/z, -fy y/z^2]]
    auto dJ = [&](int r, int col) {
      return dTm[r][0] * c2.W.m[col][0] + dTm[r][1] * c2.W.m[col][1] + dTm[r][2] * c2.W.m[col][2];
    };
    const float gJ00 = dJ(0, 0), gJ02 = dJ(0, 2), gJ11 = dJ(1, 1), gJ12 = dJ(1, 2);
    // J(t): the clamped x/z, y/z (forward: limx, limy) pass no gradient to t.x, t.y where the clamp was active
    const bool x_free = !(c2.txtz < -c2.limx || c2.txtz > c2.limx), y_free = !(c2.tytz < -c2.limy || c2.tytz > c2.limy);
    const float iz = 1.f / c2.tz, iz2 = iz * iz, iz3 = iz2 * iz;
    const float dL_dtx = x_free ? -h_x * iz2 * gJ02 : 0.f;
    const float dL_dty = y_free ? -h_y * iz2 * gJ12 : 0.f;
    const float dL_dtz = -(h_x * gJ00 + h_y * gJ11) * iz2 + 2.f * (h_x * c2.tx * gJ02 + h_y * c2.ty * gJ12) * iz3;
    // t = view * mean (rotation part): mean gets view^T dL/dt
    dmean.x = s_view[0] * dL_dtx + s_view[1] * dL_dty + s_view[2] * dL_dtz;
    dmean.y = s_view[4] * dL_dtx + s_view[5] * dL_dty + s_view[6] * dL_dtz;
    dmean.z = s_view[8] * dL_dtx + s_view[9] * dL_dty + s_view[10] * dL_dtz;

    // ---------------- preprocessCUDA backward (backward.cu:346-396) ----------------
    const float* proj = s_proj;
    const float m_hw = proj[3] * mx + proj[7] * my + proj[11] * mz + proj[15];
    const float m_w = 1.0f / (m_hw + 0.0000001f);
    const float mul1 = (proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12]) * m_w * m_w;
    const float mul2 = (proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13]) * m_w * m_w;
    const float gx_ = s_m2d[3 * tid], gy_ = s_m2d[3 * tid + 1];
    dmean.x += (proj[0] * m_w - proj[3] * mul1) * gx_ + (proj[1] * m_w - proj[3] * mul2) * gy_;
    dmean.y += (proj[4] * m_w - proj[7] * mul1) * gx_ + (proj[5] * m_w - proj[7] * mul2) * gy_;
    dmean.z += (proj[8] * m_w - proj[11] * mul1) * gx_ + (proj[9] * m_w - proj[11] * mul2) * gy_;
  }
  // dL_dcov3D: coalesced write-out (zeros for Gaussians with radius 0)
  if (dL_dcov3D) {
    float2* w2 = wcov;
    w2[lane * 3] = make_float2(dcov[0], dcov[1]);
    w2[lane * 3 + 1] = make_float2(dcov[2], dcov[3]);
    w2[lane * 3 + 2] = make_float2(dcov[4], dcov[5]);
    __syncwarp();
    float2* dst = reinterpret_cast<float2*>(dL_dcov3D + wfirst * 6);
#pragma unroll
    for (int it = 0; it < 3; it++) {
      const int c = it * 32 + lane;
      if ((valid_mask >> (c / 3)) & 1u) dst[c] = w2[c];
    }
    __syncwarp();
  }
  // ---------------- SH colour backward (backward.cu:20-139) ----------------
  if (do_sh) {
    if (SH16) {
      cp_async_wait<0>();
      __syncwarp();
      if (vis) {
        const uint8_t cl = clamped[idx];
        const float g[3] = {(cl & 1) ? 0.f : s_dcol[3 * tid], (cl & 2) ? 0.f : s_dcol[3 * tid + 1],
                            (cl & 4) ? 0.f : s_dcol[3 * tid + 2]};
        const float3 dir = make_float3(mx - s_cam[0], my - s_cam[1], mz - s_cam[2]);
        float v[48];
#pragma unroll
        for (int i = 0; i < SH_ROW; i++) {
          const float4 t = wsm[row_slot<SH_ROW>(lane, i)];
          v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
        }
        const float3 dm = sh_dir_backward(D, v, dir, g);
        dmean.x += dm.x; dmean.y += dm.y; dmean.z += dm.z;
        if (dL_dcolor_masked) { s_dcol[3 * tid] = g[0]; s_dcol[3 * tid + 1] = g[1]; s_dcol[3 * tid + 2] = g[2]; }
        const float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
        float bs[16];
        if (write_sh) sh_basis(D, dir.x / len, dir.y / len, dir.z / len, bs);
#pragma unroll
        for (int i = 0; write_sh && i < SH_ROW; i++) {
          float4 o;
          o.x = bs[(4 * i) / 3] * g[(4 * i) % 3];
          o.y = bs[(4 * i + 1) / 3] * g[(4 * i + 1) % 3];
          o.z = bs[(4 * i + 2) / 3] * g[(4 * i + 2) % 3];
          o.w = bs[(4 * i + 3) / 3] * g[(4 * i + 3) % 3];
          wsm[row_slot<SH_ROW>(lane, i)] = o;
        }
      }
      __syncwarp();
      if (write_sh)
        warp_store_rows<SH_ROW>(reinterpret_cast<float4*>(dL_dsh) + wfirst * SH_ROW, wsm, vis_mask, valid_mask, true,
                                lane);
      __syncwarp();
    } else if (valid) {
      float* dsh = write_sh ? dL_dsh + (size_t)idx * M * 3 : nullptr;
      if (vis) {
        const uint8_t cl = clamped[idx];
        float3 g;
        g.x = (cl & 1) ? 0.f : s_dcol[3 * tid];
        g.y = (cl & 2) ? 0.f : s_dcol[3 * tid + 1];
        g.z = (cl & 4) ? 0.f : s_dcol[3 * tid + 2];
        const float3 dir = make_float3(mx - s_cam[0], my - s_cam[1], mz - s_cam[2]);
        if (dL_dcolor_masked) { s_dcol[3 * tid] = g.x; s_dcol[3 * tid + 1] = g.y; s_dcol[3 * tid + 2] = g.z; }
        const float3 dm = write_sh ? sh_backward(D, M, shs + (size_t)idx * M * 3, dir, g, dsh)
                                   : sh_backward_t<false>(D, M, shs + (size_t)idx * M * 3, dir, g, nullptr);
        dmean.x += dm.x; dmean.y += dm.y; dmean.z += dm.z;
      } else if (write_sh) {
        for (int i = 0; i < M * 3; i++) dsh[i] = 0.f;
      }
    }
  } else if (dL_dsh != nullptr && valid && M > 0 && !vis) {
    // (the caller zero-fills dL_dsh when the SH backward does not run at all; nothing to do here)
  }
  // ---------------- computeCov3D backward (backward.cu:278-341) ----------------
  float dscale[3] = {0.f, 0.f, 0.f};
  float4 dq = make_float4(0.f, 0.f, 0.f, 0.f);
  if (do_scale && vis) {
    const float4 q = __ldg(reinterpret_cast<const float4*>(rotations) + idx);
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    const Mat3 R = quat_to_mat(r, x, y, z);
    const float s[3] = {scale_modifier * s_scl[3 * tid], scale_modifier * s_scl[3 * tid + 1],
                        scale_modifier * s_scl[3 * tid + 2]};
    Mat3 M2, dSig;
#pragma unroll
    for (int cc = 0; cc < 3; cc++)
#pragma unroll
      for (int rr = 0; rr < 3; rr++) M2.m[cc][rr] = 2.0f * (s[rr] * R.m[cc][rr]);
    dSig.m[0][0] = dcov[0]; dSig.m[0][1] = 0.5f * dcov[1]; dSig.m[0][2] = 0.5f * dcov[2];
    dSig.m[1][0] = 0.5f * dcov[1]; dSig.m[1][1] = dcov[3]; dSig.m[1][2] = 0.5f * dcov[4];
    dSig.m[2][0] = 0.5f * dcov[2]; dSig.m[2][1] = 0.5f * dcov[4]; dSig.m[2][2] = dcov[5];
    const Mat3 dL_dM = mat3_mul(M2, dSig);
    const Mat3 Rt = mat3_transpose(R);
    Mat3 dMt = mat3_transpose(dL_dM);
#pragma unroll
    for (int k = 0; k < 3; k++)
      dscale[k] = Rt.m[k][0] * dMt.m[k][0] + Rt.m[k][1] * dMt.m[k][1] + Rt.m[k][2] * dMt.m[k][2];
#pragma unroll
    for (int k = 0; k < 3; k++)
#pragma unroll
      for (int rr = 0; rr < 3; rr++) dMt.m[k][rr] *= s[k];
    dq.x = 2 * z * (dMt.m[0][1] - dMt.m[1][0]) + 2 * y * (dMt.m[2][0] - dMt.m[0][2]) + 2 * x * (dMt.m[1][2] - dMt.m[2][1]);
    dq.y = 2 * y * (dMt.m[1][0] + dMt.m[0][1]) + 2 * z * (dMt.m[2][0] + dMt.m[0][2]) + 2 * r * (dMt.m[1][2] - dMt.m[2][1]) -
           4 * x * (dMt.m[2][2] + dMt.m[1][1]);
    dq.z = 2 * x * (dMt.m[1][0] + dMt.m[0][1]) + 2 * r * (dMt.m[2][0] - dMt.m[0][2]) + 2 * z * (dMt.m[1][2] + dMt.m[2][1]) -
           4 * y * (dMt.m[2][2] + dMt.m[0][0]);
    dq.w = 2 * r * (dMt.m[0][1] - dMt.m[1][0]) + 2 * x * (dMt.m[2][0] + dMt.m[0][2]) + 2 * y * (dMt.m[1][2] + dMt.m[2][1]) -
           4 * z * (dMt.m[1][1] + dMt.m[0][0]);
  }
  if (do_sh && dL_dcolor_masked && !vis) { s_dcol[3 * tid] = 0.f; s_dcol[3 * tid + 1] = 0.f; s_dcol[3 * tid + 2] = 0.f; }
  if (do_scale && valid) reinterpret_cast<float4*>(dL_drot)[idx] = dq;
  // dL_dmean3D / dL_dscale leave through the arrays their inputs came in by (each thread owns its triplet)
  __syncthreads();
  s_xyz[3 * tid] = dmean.x; s_xyz[3 * tid + 1] = dmean.y; s_xyz[3 * tid + 2] = dmean.z;
  s_scl[3 * tid] = dscale[0]; s_scl[3 * tid + 1] = dscale[1]; s_scl[3 * tid + 2] = dscale[2];
  __syncthreads();
  if (dL_dmean3D) unstage_flat<BPRE_THREADS>(dL_dmean3D, s_xyz, (size_t)first * 3, count * 3, tid);
  if (moments || acc) unstage_flat<BPRE_THREADS>(dL_dmean2D, s_m2d, (size_t)first * 3, count * 3, tid);
  if (do_sh && dL_dcolor_masked) unstage_flat<BPRE_THREADS>(dL_dcolor_masked, s_dcol, (size_t)first * 3, count * 3, tid);
  if (do_scale) unstage_flat<BPRE_THREADS>(dL_dscale, s_scl, (size_t)first * 3, count * 3, tid);
}

// ---------------------------------------------------------------------------------------------------------------
// dL_dsh from per-view factors.  For one view, dL_dsh[k][ch] = basis_k(normalize(mean - campos)) * g[ch]
// (backward.cu:36-129), a rank-1 matrix per Gaussian, g = that view's clamp-masked dL_dcolor (zero if culled).  Given
// the factors g_v of V views (the local ones or those all-gathered from the other ranks: 12 B per Gaussian and view
// instead of 192), this kernel forms sum_v basis(dir_v) (x) g_v and writes the [P,M,3] gradient once, rows staged
// through shared memory for coalesced 16-byte stores when M == 16.
// ---------------------------------------------------------------------------------------------------------------
constexpr int SHV_THREADS = 128;
constexpr int SHV_BATCH = 8;  // views whose factors are loaded together (independent loads in flight: NVLink latency)
constexpr size_t SHV_SMEM = (size_t)(SHV_THREADS / 32) * 32 * (SH_ROW + 1) * sizeof(float4);

// Per-view base pointers: g[v] -> that view's [P,3] factor, campos[v] -> its 3 floats.  The pointers may be peer
// memory (another GPU's buffer mapped over NVLink, lsr_peer_open): the kernel then IS the exchange — every rank reads
// the other ranks' factors in place, no all-gather staging copy.
struct ShViewPtrs {
  const float* g[LSR_MAX_VIEWS];
  const float* campos[LSR_MAX_VIEWS];
};

template <bool SH16>
__global__ void __launch_bounds__(SHV_THREADS)
sh_grad_views_kernel(int P, int D, int M, int V, const float* __restrict__ means3D, const ShViewPtrs vp,
                     float* __restrict__ dL_dsh, int accumulate) {
  extern __shared__ __align__(16) float4 s_rows[];
  __shared__ float s_cam[LSR_MAX_VIEWS * 3];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int idx = blockIdx.x * SHV_THREADS + tid;
  const bool valid = idx < P;
  for (int i = tid; i < V * 3; i += SHV_THREADS) s_cam[i] = vp.campos[i / 3][i % 3];
  float acc[48];
#pragma unroll
  for (int i = 0; i < 48; i++) acc[i] = 0.f;
  float mx = 0.f, my = 0.f, mz = 0.f;
  if (valid) { mx = __ldg(means3D + 3 * (size_t)idx); my = __ldg(means3D + 3 * (size_t)idx + 1); mz = __ldg(means3D + 3 * (size_t)idx + 2); }
  __syncthreads();
  for (int v0 = 0; v0 < V; v0 += SHV_BATCH) {
    float gv[SHV_BATCH][3];
#pragma unroll
    for (int j = 0; j < SHV_BATCH; j++) {
      gv[j][0] = gv[j][1] = gv[j][2] = 0.f;
      if (valid && v0 + j < V) {
        const float* gp = vp.g[v0 + j] + (size_t)idx * 3;
        gv[j][0] = gp[0]; gv[j][1] = gp[1]; gv[j][2] = gp[2];
      }
    }
#pragma unroll 1
    for (int j = 0; j < SHV_BATCH; j++) {
      const float g0 = gv[j][0], g1 = gv[j][1], g2 = gv[j][2];
      if (g0 == 0.f && g1 == 0.f && g2 == 0.f) continue;  // culled in this view (or no gradient): contributes nothing
      const int v = v0 + j;
      const float dx = mx - s_cam[3 * v], dy = my - s_cam[3 * v + 1], dz = mz - s_cam[3 * v + 2];
      const float len = sqrtf(dx * dx + dy * dy + dz * dz);
      float bs[16];
      sh_basis(D, dx / len, dy / len, dz / len, bs);
#pragma unroll
      for (int k = 0; k < 16; k++) {
        acc[3 * k] += bs[k] * g0;
        acc[3 * k + 1] += bs[k] * g1;
        acc[3 * k + 2] += bs[k] * g2;
      }
    }
  }
  if (SH16) {
    float4* wsm = s_rows + (size_t)warp * 32 * (SH_ROW + 1);
#pragma unroll
    for (int i = 0; i < SH_ROW; i++)
      wsm[row_slot<SH_ROW>(lane, i)] = make_float4(acc[4 * i], acc[4 * i + 1], acc[4 * i + 2], acc[4 * i + 3]);
    __syncwarp();
    const uint32_t valid_mask = __ballot_sync(0xffffffffu, valid);
    warp_store_rows<SH_ROW>(reinterpret_cast<float4*>(dL_dsh) + (size_t)(blockIdx.x * SHV_THREADS + warp * 32) * SH_ROW,
                            wsm, valid_mask, valid_mask, false, lane);
  } else if (valid) {
    float* o = dL_dsh + (size_t)idx * M * 3;
    if (accumulate) { for (int i = 0; i < M * 3 && i < 48; i++) o[i] += acc[i]; }
    else { for (int i = 0; i < M * 3; i++) o[i] = i < 48 ? acc[i] : 0.f; }
  }
}

// acc -> the per-tensor gradients, for backward passes that do not run the per-Gaussian kernel (feature-only training)
__global__ void __launch_bounds__(256)
unpack_acc_kernel(int P, const float4* __restrict__ acc, float* __restrict__ dL_dmean2D, float* __restrict__ dL_dcolors,
                  float* __restrict__ dL_dfeat) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  if (dL_dmean2D) {
    const float4 m = __ldg(acc + 4 * (size_t)idx);
    dL_dmean2D[3 * (size_t)idx] = m.x; dL_dmean2D[3 * (size_t)idx + 1] = m.y; dL_dmean2D[3 * (size_t)idx + 2] = 0.f;
  }
  if (dL_dcolors) {
    const float4 c = __ldg(acc + 4 * (size_t)idx + 2);
    dL_dcolors[3 * (size_t)idx] = c.x; dL_dcolors[3 * (size_t)idx + 1] = c.y; dL_dcolors[3 * (size_t)idx + 2] = c.z;
  }
  if (dL_dfeat) {
    const float4 f = __ldg(acc + 4 * (size_t)idx + 3);
    dL_dfeat[3 * (size_t)idx] = f.x; dL_dfeat[3 * (size_t)idx + 1] = f.y; dL_dfeat[3 * (size_t)idx + 2] = f.z;
  }
}

}  // namespace

void launch_unpack_acc(int P, const float* acc, float* dL_dmean2D, float* dL_dcolors, float* dL_dfeat, cudaStream_t s) {
  if (P <= 0) return;
  unpack_acc_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, reinterpret_cast<const float4*>(acc), dL_dmean2D, dL_dcolors,
                                                    dL_dfeat);
  count_launch();
}

void launch_preprocess_bwd(const lsr_frame& f, const lsr_inputs& in, float focal_x, float focal_y,
                           const int32_t* radii, GeomView g, uint32_t need_mask, const lsr_grads& gr, const float* acc,
                           cudaStream_t s) {
  if (f.P <= 0) return;
  // the blend kernel leaves moments instead of gradients exactly when it ran in its GEOM mode (render_bwd.cu)
  const int moments = (LSR_MOMENTS && (need_mask & LSR_NEED_GEOMETRY)) ? 1 : 0;
  const float* cov3Ds = in.cov3D_precomp ? in.cov3D_precomp : g.cov3D;
  const bool sh16 = in.shs != nullptr && (gr.dL_dsh != nullptr || gr.dL_dcolors_masked != nullptr) && f.M == 16 &&
                    (((uintptr_t)in.shs) & 15) == 0 && (((uintptr_t)gr.dL_dsh) & 15) == 0;
  static DeviceOnce attr;  // per device: a process may drive several GPUs
  attr.run([] {
    cudaFuncSetAttribute(preprocess_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BPRE_SH_SMEM);
    cudaFuncSetAttribute(preprocess_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BPRE_SH_SMEM);
  });
  const int blocks = (f.P + BPRE_THREADS - 1) / BPRE_THREADS;
#define LSR_BPRE_ARGS                                                                                              \
  f.P, f.D, f.M, in.means3D, radii, in.shs, g.clamped, in.scales, in.rotations, f.scale_modifier, cov3Ds,          \
      f.viewmatrix, f.projmatrix, focal_x, focal_y, f.tanfovx, f.tanfovy, f.campos, in.opacities, moments,       \
      0.5f * f.width,                                                                                              \
      0.5f * f.height, gr.dL_dmeans2D, gr.dL_dconic,                                                               \
      gr.dL_dcolors, gr.dL_dmeans3D, gr.dL_dcov3D, gr.dL_dsh, gr.dL_dscales, gr.dL_drotations,                      \
      gr.dL_dcolors_masked, reinterpret_cast<const float4*>(acc), acc ? gr.dL_dopacity : nullptr,                    \
      acc ? gr.dL_dcolors : nullptr, acc ? gr.dL_dlanguage_feature : nullptr
  if (sh16) preprocess_bwd_kernel<true><<<blocks, BPRE_THREADS, BPRE_SH_SMEM, s>>>(LSR_BPRE_ARGS);
  else preprocess_bwd_kernel<false><<<blocks, BPRE_THREADS, BPRE_SH_SMEM, s>>>(LSR_BPRE_ARGS);
#undef LSR_BPRE_ARGS
  count_launch();
}

void launch_sh_grad_view_ptrs(int P, int D, int M, int V, const float* means3D, const float* const* g_views,
                              const float* const* campos_views, float* dL_dsh, cudaStream_t s) {
  if (P <= 0 || M <= 0) return;
  const int blocks = (P + SHV_THREADS - 1) / SHV_THREADS;
  // LSR_MAX_VIEWS views per launch; later chunks add to what the first one wrote (scalar path)
  for (int v0 = 0; v0 == 0 || v0 < V; v0 += LSR_MAX_VIEWS) {
    const int n = std::min(V - v0, (int)LSR_MAX_VIEWS);
    ShViewPtrs vp;
    for (int i = 0; i < LSR_MAX_VIEWS; i++) {
      vp.g[i] = i < n ? g_views[v0 + i] : nullptr;
      vp.campos[i] = i < n ? campos_views[v0 + i] : nullptr;
    }
    const bool sh16 = v0 == 0 && M == 16 && (((uintptr_t)dL_dsh) & 15) == 0;
    if (sh16) sh_grad_views_kernel<true><<<blocks, SHV_THREADS, SHV_SMEM, s>>>(P, D, M, n, means3D, vp, dL_dsh, 0);
    else sh_grad_views_kernel<false><<<blocks, SHV_THREADS, SHV_SMEM, s>>>(P, D, M, n, means3D, vp, dL_dsh, v0 > 0);
    count_launch();
  }
}

void launch_sh_grad_views(int P, int D, int M, int V, const float* means3D, const float* campos, const float* g,
                          float* dL_dsh, cudaStream_t s) {
  std::vector<const float*> gp((size_t)std::max(V, 0)), cp((size_t)std::max(V, 0));
  for (int v = 0; v < V; v++) { gp[v] = g + (size_t)v * P * 3; cp[v] = campos + 3 * (size_t)v; }
  launch_sh_grad_view_ptrs(P, D, M, V, means3D, gp.data(), cp.data(), dL_dsh, s);
}

}  // namespace lsr
