// adam.cu — the optimizer step that follows the backward, fused over all parameter groups.
// The reference builds torch.optim.Adam(l, lr=0.0, eps=1e-15) over six groups (xyz, f_dc, f_rest, opacity, scaling,
// rotation) or one (language_feature) and calls optimizer.step() every iteration (scene/gaussian_model.py:198-227,
// train.py:136-138): about ten elementwise passes per group, each streaming the [P,·] tensors through HBM.  Here: ONE
// launch for all groups, one pass — 16 B read (param, grad, exp_avg, exp_avg_sq) and 12 B written per element, i.e.
// 1652 B per Gaussian for the 59 RGB-mode parameters.  HBM-bound; float4 accesses, 4 x 4 elements per thread.
// Arithmetic is PyTorch's single-tensor Adam (torch/optim/adam.py, amsgrad=False, weight_decay=0, maximize=False):
//     exp_avg.lerp_(grad, 1 - beta1);  exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
//     denom = exp_avg_sq.sqrt() / sqrt(1 - beta2^t) + eps;  param.addcdiv_(exp_avg, denom, value=-lr / (1 - beta1^t))
// with the step-dependent scalars formed in double on the host like Python does.
// Optional `visible` [P] (uint8): only the rows of visible Gaussians are touched — moments included — which is the
// "sparse Adam" of later 3DGS releases; null reproduces the reference (dense).
#include <cmath>

#include "common.cuh"

namespace lsr {

namespace {

constexpr int ADAM_THREADS = 256;
constexpr int ADAM_VEC = 4;                                   // float4 per thread
constexpr int ADAM_BLOCK_ELEMS = ADAM_THREADS * ADAM_VEC * 4;  // 4096 elements per CTA

struct AdamArgs {
  float* param[LSR_ADAM_MAX_TENSORS];
  const float* grad[LSR_ADAM_MAX_TENSORS];
  float* m[LSR_ADAM_MAX_TENSORS];
  float* v[LSR_ADAM_MAX_TENSORS];
  int64_t numel[LSR_ADAM_MAX_TENSORS];
  int32_t row[LSR_ADAM_MAX_TENSORS];         // elements per Gaussian (for the visibility mask)
  float neg_step_size[LSR_ADAM_MAX_TENSORS];  // -lr / (1 - beta1^t)
  uint32_t block_start[LSR_ADAM_MAX_TENSORS + 1];
  int n;
};

__device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, float w1, float beta2, float w2,
                                            float inv_bc2_sqrt, float eps, float neg_step) {
  m = m + w1 * (g - m);                      // lerp, |weight| < 0.5 branch
  v = v * beta2;
  v = v + (w2 * g) * g;                      // addcmul: a + alpha * b * c
  const float denom = __fsqrt_rn(v) * inv_bc2_sqrt + eps;
  p = p + neg_step * __fdiv_rn(m, denom);    // addcdiv: a + alpha * (b / c)
}

__global__ void __launch_bounds__(ADAM_THREADS)
adam_kernel(const AdamArgs a, float w1, float beta2, float w2, float inv_bc2_sqrt, float eps,
            const uint8_t* __restrict__ visible) {
  int t = 0;
  while (t + 1 < a.n && blockIdx.x >= a.block_start[t + 1]) t++;
  const int64_t base = (int64_t)(blockIdx.x - a.block_start[t]) * ADAM_BLOCK_ELEMS;
  const int64_t n = a.numel[t];
  float* __restrict__ P = a.param[t];
  const float* __restrict__ G = a.grad[t];
  float* __restrict__ M = a.m[t];
  float* __restrict__ V = a.v[t];
  const float neg_step = a.neg_step_size[t];
  const int row = a.row[t];
  const bool vec = (((uintptr_t)P | (uintptr_t)G | (uintptr_t)M | (uintptr_t)V) & 15) == 0;
#pragma unroll
  for (int it = 0; it < 4; it++) {
    const int64_t e = base + ((int64_t)it * ADAM_THREADS + threadIdx.x) * ADAM_VEC;
    if (e >= n) break;
    if (vec && e + ADAM_VEC <= n) {
      float4 p = *reinterpret_cast<const float4*>(P + e);
      const float4 g = __ldg(reinterpret_cast<const float4*>(G + e));
      float4 m = *reinterpret_cast<const float4*>(M + e);
      float4 v = *reinterpret_cast<const float4*>(V + e);
      bool on[4] = {true, true, true, true};
      if (visible) {
#pragma unroll
        for (int k = 0; k < 4; k++) on[k] = visible[(e + k) / row] != 0;
        if (!(on[0] | on[1] | on[2] | on[3])) continue;
      }
      if (on[0]) adam_update(p.x, g.x, m.x, v.x, w1, beta2, w2, inv_bc2_sqrt, eps, neg_step);
      if (on[1]) adam_update(p.y, g.y, m.y, v.y, w1, beta2, w2, inv_bc2_sqrt, eps, neg_step);
      if (on[2]) adam_update(p.z, g.z, m.z, v.z, w1, beta2, w2, inv_bc2_sqrt, eps, neg_step);
      if (on[3]) adam_update(p.w, g.w, m.w, v.w, w1, beta2, w2, inv_bc2_sqrt, eps, neg_step);
      *reinterpret_cast<float4*>(P + e) = p;
      *reinterpret_cast<float4*>(M + e) = m;
      *reinterpret_cast<float4*>(V + e) = v;
    } else {
      for (int k = 0; k < ADAM_VEC && e + k < n; k++) {
        if (visible && visible[(e + k) / row] == 0) continue;
        float p = P[e + k], m = M[e + k], v = V[e + k];
        adam_update(p, __ldg(G + e + k), m, v, w1, beta2, w2, inv_bc2_sqrt, eps, neg_step);
        P[e + k] = p; M[e + k] = m; V[e + k] = v;
      }
    }
  }
}

}  // namespace

void launch_adam(int n, const lsr_adam_tensor* ts, int64_t step, double beta1, double beta2, double eps,
                 const uint8_t* visible, cudaStream_t s) {
  AdamArgs a{};
  a.n = 0;
  uint32_t blocks = 0;
  const double bc1 = 1.0 - std::pow(beta1, (double)step), bc2 = 1.0 - std::pow(beta2, (double)step);
  for (int i = 0; i < n; i++) {
    if (ts[i].numel <= 0) continue;
    const int k = a.n++;
    a.param[k] = ts[i].param; a.grad[k] = ts[i].grad; a.m[k] = ts[i].exp_avg; a.v[k] = ts[i].exp_avg_sq;
    a.numel[k] = ts[i].numel;
    a.row[k] = ts[i].row_width > 0 ? ts[i].row_width : 1;
    a.neg_step_size[k] = (float)(-((double)ts[i].lr / bc1));
    a.block_start[k] = blocks;
    blocks += (uint32_t)((ts[i].numel + ADAM_BLOCK_ELEMS - 1) / ADAM_BLOCK_ELEMS);
  }
  a.block_start[a.n] = blocks;
  if (blocks == 0) return;
  const float inv_bc2_sqrt = 1.0f / (float)std::sqrt(bc2);  // ATen divides by a CPU scalar as a * (1 / b) in float
  adam_kernel<<<blocks, ADAM_THREADS, 0, s>>>(a, (float)(1.0 - beta1), (float)beta2, (float)(1.0 - beta2),
                                              inv_bc2_sqrt, (float)eps, visible);
  count_launch();
}

}  // namespace lsr
