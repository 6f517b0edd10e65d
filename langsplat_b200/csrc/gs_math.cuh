// gs_math.cuh — per-Gaussian projection math with the rounding sequence pinned.
//
// radii / depth / means2D / tile rect must be bit-identical to the reference build, whose values depend on how
// nvcc contracts the GLM expression templates.  Everything on that chain is therefore
// written with explicit round-to-nearest intrinsics (__fmul_rn / __fmaf_rn / __fadd_rn / __fdiv_rn / __fsqrt_rn),
// which the compiler neither contracts nor reassociates, in the order the reference's SASS performs them:
//   sum of products  p1 + p2 + p3 (+ c)  ->  t = a2*b2; t = fma(a1,b1,t); t = fma(a3,b3,t); t = t + c.
#pragma once
#include <cuda_runtime.h>

namespace lsr {

struct Mat3 {  // GLM convention: m[c][r], column major
  float m[3][3];
};

// glm::mat3 * glm::mat3 (RAST/third_party/glm/glm/detail/type_mat3x3.inl:486-517) as nvcc contracts it.
__device__ __forceinline__ Mat3 mat3_mul(const Mat3& A, const Mat3& B) {
  Mat3 C;
#pragma unroll
  for (int c = 0; c < 3; c++)
#pragma unroll
    for (int r = 0; r < 3; r++) {
      float t = __fmul_rn(A.m[1][r], B.m[c][1]);
      t = __fmaf_rn(A.m[0][r], B.m[c][0], t);
      t = __fmaf_rn(A.m[2][r], B.m[c][2], t);
      C.m[c][r] = t;
    }
  return C;
}
__device__ __forceinline__ Mat3 mat3_transpose(const Mat3& A) {
  Mat3 C;
#pragma unroll
  for (int c = 0; c < 3; c++)
#pragma unroll
    for (int r = 0; r < 3; r++) C.m[c][r] = A.m[r][c];
  return C;
}

// one row of transformPoint4x3 / 4x4 — RAST/cuda_rasterizer/auxiliary.h:58-77
__device__ __forceinline__ float tp_row(const float* __restrict__ m, int i, float x, float y, float z) {
  float t = __fmul_rn(m[4 + i], y);
  t = __fmaf_rn(m[i], x, t);
  t = __fmaf_rn(m[8 + i], z, t);
  return __fadd_rn(t, m[12 + i]);
}

// rotation matrix from the quaternion as given (not normalised) — RAST/cuda_rasterizer/forward.cu:127-139.
// Contraction read off the reference build's SASS (nvcc 12.9, sm_100a): products with several uses stay separate
// FMULs, so  yy+zz = (y*y) + (z*z)  [no fma],  xz+ry = fma(r,y, x*z),  xz-ry = fma(-r,y, x*z);  the others fuse
// their left product:  xy-rz = fma(x,y,-(r*z)),  xx+zz = fma(x,x, z*z), ...
__device__ __forceinline__ Mat3 quat_to_mat(float r, float x, float y, float z) {
  Mat3 R;
  float t;
  t = __fadd_rn(__fmul_rn(y, y), __fmul_rn(z, z)); R.m[0][0] = __fadd_rn(1.f, -__fadd_rn(t, t));
  t = __fmaf_rn(x, y, -__fmul_rn(r, z));          R.m[0][1] = __fadd_rn(t, t);
  t = __fmaf_rn(r, y, __fmul_rn(x, z));           R.m[0][2] = __fadd_rn(t, t);
  t = __fmaf_rn(x, y, __fmul_rn(r, z));           R.m[1][0] = __fadd_rn(t, t);
  t = __fmaf_rn(x, x, __fmul_rn(z, z));           R.m[1][1] = __fadd_rn(1.f, -__fadd_rn(t, t));
  t = __fmaf_rn(y, z, -__fmul_rn(r, x));          R.m[1][2] = __fadd_rn(t, t);
  t = __fmaf_rn(-r, y, __fmul_rn(x, z));          R.m[2][0] = __fadd_rn(t, t);
  t = __fmaf_rn(y, z, __fmul_rn(r, x));           R.m[2][1] = __fadd_rn(t, t);
  t = __fmaf_rn(x, x, __fmul_rn(y, y));           R.m[2][2] = __fadd_rn(1.f, -__fadd_rn(t, t));
  return R;
}

// RAST/cuda_rasterizer/forward.cu:119-153.  M = S*R -> M[c][r] = s_r * R[c][r]; Sigma = M^T * M.
__device__ __forceinline__ void compute_cov3D(float sx, float sy, float sz, float mod, float qr, float qx, float qy,
                                              float qz, float* cov3D, Mat3* R_out = nullptr, Mat3* M_out = nullptr) {
  Mat3 R = quat_to_mat(qr, qx, qy, qz);
  const float s[3] = {__fmul_rn(mod, sx), __fmul_rn(mod, sy), __fmul_rn(mod, sz)};
  Mat3 M;
#pragma unroll
  for (int c = 0; c < 3; c++)
#pragma unroll
    for (int r = 0; r < 3; r++) M.m[c][r] = __fmul_rn(s[r], R.m[c][r]);
  Mat3 Sigma = mat3_mul(mat3_transpose(M), M);
  cov3D[0] = Sigma.m[0][0];
  cov3D[1] = Sigma.m[0][1];
  cov3D[2] = Sigma.m[0][2];
  cov3D[3] = Sigma.m[1][1];
  cov3D[4] = Sigma.m[1][2];
  cov3D[5] = Sigma.m[2][2];
  if (R_out) *R_out = R;
  if (M_out) *M_out = M;
}

struct Cov2D {
  float tx, ty, tz, txtz, tytz, limx, limy;
  Mat3 T, Vrk, W;
  float a, b, c;  // cov[0][0]+0.3, cov[0][1], cov[1][1]+0.3
};

// RAST/cuda_rasterizer/forward.cu:75-114 (re-evaluated identically by backward.cu:155-201)
__device__ __forceinline__ void compute_cov2D(float mx, float my, float mz, float fx, float fy, float tan_fovx,
                                              float tan_fovy, const float* cov3D, const float* __restrict__ V,
                                              Cov2D& o) {
  float tx = tp_row(V, 0, mx, my, mz);
  float ty = tp_row(V, 1, mx, my, mz);
  const float tz = tp_row(V, 2, mx, my, mz);
  o.limx = __fmul_rn(1.3f, tan_fovx);
  o.limy = __fmul_rn(1.3f, tan_fovy);
  o.txtz = __fdiv_rn(tx, tz);
  o.tytz = __fdiv_rn(ty, tz);
  tx = __fmul_rn(fminf(o.limx, fmaxf(-o.limx, o.txtz)), tz);
  ty = __fmul_rn(fminf(o.limy, fmaxf(-o.limy, o.tytz)), tz);
  o.tx = tx; o.ty = ty; o.tz = tz;
  Mat3 J;
  const float tz2 = __fmul_rn(tz, tz);
  J.m[0][0] = __fdiv_rn(fx, tz); J.m[0][1] = 0.f; J.m[0][2] = __fdiv_rn(-__fmul_rn(fx, tx), tz2);
  J.m[1][0] = 0.f; J.m[1][1] = __fdiv_rn(fy, tz); J.m[1][2] = __fdiv_rn(-__fmul_rn(fy, ty), tz2);
  J.m[2][0] = 0.f; J.m[2][1] = 0.f; J.m[2][2] = 0.f;
  Mat3& W = o.W;
  W.m[0][0] = V[0]; W.m[0][1] = V[4]; W.m[0][2] = V[8];
  W.m[1][0] = V[1]; W.m[1][1] = V[5]; W.m[1][2] = V[9];
  W.m[2][0] = V[2]; W.m[2][1] = V[6]; W.m[2][2] = V[10];
  o.T = mat3_mul(W, J);
  Mat3& K = o.Vrk;
  K.m[0][0] = cov3D[0]; K.m[0][1] = cov3D[1]; K.m[0][2] = cov3D[2];
  K.m[1][0] = cov3D[1]; K.m[1][1] = cov3D[3]; K.m[1][2] = cov3D[4];
  K.m[2][0] = cov3D[2]; K.m[2][1] = cov3D[4]; K.m[2][2] = cov3D[5];
  Mat3 cov = mat3_mul(mat3_mul(mat3_transpose(o.T), mat3_transpose(K)), o.T);
  o.a = __fadd_rn(cov.m[0][0], 0.3f);
  o.b = cov.m[0][1];
  o.c = __fadd_rn(cov.m[1][1], 0.3f);
}

// RAST/cuda_rasterizer/auxiliary.h:41-44 — evaluated in double (DADD, DFMA, DMUL in the reference SASS)
__device__ __forceinline__ float ndc2pix(float v, int S) {
  double t = __dadd_rn((double)v, 1.0);
  t = __fma_rn(t, (double)S, -1.0);
  return (float)__dmul_rn(t, 0.5);
}

// RAST/cuda_rasterizer/auxiliary.h:46-56
__device__ __forceinline__ void get_rect(float px, float py, int radius, int gx, int gy, uint32_t& x0, uint32_t& y0,
                                         uint32_t& x1, uint32_t& y1) {
  const float r = (float)radius;
  x0 = min((uint32_t)gx, (uint32_t)max(0, __float2int_rz(__fmul_rn(__fadd_rn(px, -r), 0.0625f))));
  y0 = min((uint32_t)gy, (uint32_t)max(0, __float2int_rz(__fmul_rn(__fadd_rn(py, -r), 0.0625f))));
  x1 = min((uint32_t)gx,
           (uint32_t)max(0, __float2int_rz(__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(px, r), 16.f), -1.f), 0.0625f))));
  y1 = min((uint32_t)gy,
           (uint32_t)max(0, __float2int_rz(__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(py, r), 16.f), -1.f), 0.0625f))));
}

// SH basis constants — RAST/cuda_rasterizer/auxiliary.h:22-39
#define LSR_SH_C0 0.28209479177387814f
#define LSR_SH_C1 0.4886025119029199f
#define LSR_SH_C2_0 1.0925484305920792f
#define LSR_SH_C2_1 -1.0925484305920792f
#define LSR_SH_C2_2 0.31539156525252005f
#define LSR_SH_C2_3 -1.0925484305920792f
#define LSR_SH_C2_4 0.5462742152960396f
#define LSR_SH_C3_0 -0.5900435899266435f
#define LSR_SH_C3_1 2.890611442640554f
#define LSR_SH_C3_2 -0.4570457994644658f
#define LSR_SH_C3_3 0.3731763325901154f
#define LSR_SH_C3_4 -0.4570457994644658f
#define LSR_SH_C3_5 1.445305721320277f
#define LSR_SH_C3_6 -0.5900435899266435f

// SH basis values b[k]: RGB[ch] = sum_k b[k] * sh[k][ch] + 0.5 (forward.cu:21-72) and dRGB/dsh[k][ch] = b[k]
// (backward.cu:36-129); zero above the active degree.
__device__ __forceinline__ void sh_basis(int deg, float x, float y, float z, float (&b)[16]) {
#pragma unroll
  for (int k = 0; k < 16; k++) b[k] = 0.f;
  b[0] = LSR_SH_C0;
  if (deg > 0) {
    b[1] = -LSR_SH_C1 * y; b[2] = LSR_SH_C1 * z; b[3] = -LSR_SH_C1 * x;
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      b[4] = LSR_SH_C2_0 * xy; b[5] = LSR_SH_C2_1 * yz; b[6] = LSR_SH_C2_2 * (2.f * zz - xx - yy);
      b[7] = LSR_SH_C2_3 * xz; b[8] = LSR_SH_C2_4 * (xx - yy);
      if (deg > 2) {
        b[9] = LSR_SH_C3_0 * y * (3.f * xx - yy); b[10] = LSR_SH_C3_1 * xy * z;
        b[11] = LSR_SH_C3_2 * y * (4.f * zz - xx - yy); b[12] = LSR_SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy);
        b[13] = LSR_SH_C3_4 * x * (4.f * zz - xx - yy); b[14] = LSR_SH_C3_5 * z * (xx - yy);
        b[15] = LSR_SH_C3_6 * x * (xx - 3.f * yy);
      }
    }
  }
}


}  // namespace lsr
