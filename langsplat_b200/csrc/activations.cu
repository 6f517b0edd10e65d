// activations.cu — the per-Gaussian activations LangSplat applies right before the rasterizer, fused (SURVEY.md 8(f)
// row 1, the half in front of the path).  Per frame the reference runs, each as its own PyTorch kernel(s) with a [P,·]
// temporary,
//     scales    = torch.exp(_scaling)                       scene/gaussian_model.py:33,134-136
//     rotations = torch.nn.functional.normalize(_rotation)  scene/gaussian_model.py:40,138-140  (x / max(||x||, 1e-12))
//     opacity   = torch.sigmoid(_opacity)                   scene/gaussian_model.py:37,153-155
//     language_feature_precomp = f / (f.norm(dim=-1, keepdim=True) + 1e-9)   gaussian_renderer/__init__.py:86-88
// and autograd replays about twice as many backward.  Here: one forward pass (44 B read + 44 B written per Gaussian)
// and one backward pass that recomputes the activations from the raw parameters instead of saving them (44 + 44 B
// read, 44 B written).  HBM-bound elementwise work, one thread per Gaussian; the 12-byte rows of a warp are
// contiguous, so every 32-byte sector fetched is fully used.
// Arithmetic: expf / IEEE divide / sqrt as PyTorch's kernels use them; the two norms are summed left to right, which
// may differ from PyTorch's reduction order in the last bit (tests hold 2e-7 relative, bit-exact for exp / sigmoid).
#include "common.cuh"

namespace lsr {

namespace {

constexpr int ACT_THREADS = 256;
constexpr float ROT_EPS = 1e-12f;   // F.normalize default eps
constexpr float FEAT_EPS = 1e-9f;   // gaussian_renderer/__init__.py:87

__device__ __forceinline__ float sigmoidf(float x) {
  return __fdiv_rn(1.f, __fadd_rn(1.f, expf(-x)));  // ATen sigmoid: 1 / (1 + exp(-x))
}

__global__ void __launch_bounds__(ACT_THREADS)
activate_forward_kernel(int P, const float* __restrict__ scaling, const float* __restrict__ rotation,
                        const float* __restrict__ opacity, const float* __restrict__ feature,
                        float* __restrict__ scales_out, float* __restrict__ rotations_out,
                        float* __restrict__ opacities_out, float* __restrict__ feature_out) {
  const int i = blockIdx.x * ACT_THREADS + threadIdx.x;
  if (i >= P) return;
  if (scaling) {
    const float* s = scaling + 3 * (size_t)i;
    float* o = scales_out + 3 * (size_t)i;
    const float a = __ldg(s), b = __ldg(s + 1), c = __ldg(s + 2);
    o[0] = expf(a); o[1] = expf(b); o[2] = expf(c);
  }
  if (rotation) {
    const float4 q = __ldg(reinterpret_cast<const float4*>(rotation) + i);
    const float n = __fsqrt_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(q.x, q.x), __fmul_rn(q.y, q.y)),
                                                   __fmul_rn(q.z, q.z)), __fmul_rn(q.w, q.w)));
    const float d = fmaxf(n, ROT_EPS);
    reinterpret_cast<float4*>(rotations_out)[i] =
        make_float4(__fdiv_rn(q.x, d), __fdiv_rn(q.y, d), __fdiv_rn(q.z, d), __fdiv_rn(q.w, d));
  }
  if (opacity) opacities_out[i] = sigmoidf(__ldg(opacity + i));
  if (feature) {
    const float* f = feature + 3 * (size_t)i;
    float* o = feature_out + 3 * (size_t)i;
    const float a = __ldg(f), b = __ldg(f + 1), c = __ldg(f + 2);
    const float n = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b)), __fmul_rn(c, c)));
    const float d = __fadd_rn(n, FEAT_EPS);
    o[0] = __fdiv_rn(a, d); o[1] = __fdiv_rn(b, d); o[2] = __fdiv_rn(c, d);
  }
}

// Backward of the four activations; every gradient pair (upstream, output) is optional.
//   exp:       dx = dy * y
//   normalize: y = x / d, d = max(n, eps):  dx = dy / d - [n >= eps] * (x / n) * sum(dy * x) / d^2
//   sigmoid:   dx = dy * y * (1 - y)
//   feature:   y = f / (n + eps):           dx = dy / d - [n > 0] * (f / n) * sum(dy * f) / d^2
// (autograd of the reference expressions: clamp_min passes the gradient where n >= eps; norm's backward is 0 at n = 0)
__global__ void __launch_bounds__(ACT_THREADS)
activate_backward_kernel(int P, const float* __restrict__ scaling, const float* __restrict__ rotation,
                         const float* __restrict__ opacity, const float* __restrict__ feature,
                         const float* __restrict__ g_scales, const float* __restrict__ g_rotations,
                         const float* __restrict__ g_opacities, const float* __restrict__ g_feature,
                         float* __restrict__ d_scaling, float* __restrict__ d_rotation,
                         float* __restrict__ d_opacity, float* __restrict__ d_feature) {
  const int i = blockIdx.x * ACT_THREADS + threadIdx.x;
  if (i >= P) return;
  if (d_scaling) {
    const float* s = scaling + 3 * (size_t)i;
    const float* g = g_scales + 3 * (size_t)i;
    float* o = d_scaling + 3 * (size_t)i;
    const float a = __ldg(s), b = __ldg(s + 1), c = __ldg(s + 2);
    const float ga = __ldg(g), gb = __ldg(g + 1), gc = __ldg(g + 2);
    o[0] = ga * expf(a); o[1] = gb * expf(b); o[2] = gc * expf(c);
  }
  if (d_rotation) {
    const float4 q = __ldg(reinterpret_cast<const float4*>(rotation) + i);
    const float4 g = __ldg(reinterpret_cast<const float4*>(g_rotations) + i);
    const float n = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    const float d = fmaxf(n, ROT_EPS);
    const float inv = 1.f / d;
    float k = 0.f;
    if (n >= ROT_EPS) k = (g.x * q.x + g.y * q.y + g.z * q.z + g.w * q.w) * inv * inv / n;
    reinterpret_cast<float4*>(d_rotation)[i] =
        make_float4(g.x * inv - q.x * k, g.y * inv - q.y * k, g.z * inv - q.z * k, g.w * inv - q.w * k);
  }
  if (d_opacity) {
    const float y = sigmoidf(__ldg(opacity + i));
    d_opacity[i] = __ldg(g_opacities + i) * ((1.f - y) * y);
  }
  if (d_feature) {
    const float* f = feature + 3 * (size_t)i;
    const float* g = g_feature + 3 * (size_t)i;
    float* o = d_feature + 3 * (size_t)i;
    const float a = __ldg(f), b = __ldg(f + 1), c = __ldg(f + 2);
    const float ga = __ldg(g), gb = __ldg(g + 1), gc = __ldg(g + 2);
    const float n = sqrtf(a * a + b * b + c * c);
    const float inv = 1.f / (n + FEAT_EPS);
    float k = 0.f;
    if (n > 0.f) k = (ga * a + gb * b + gc * c) * inv * inv / n;
    o[0] = ga * inv - a * k; o[1] = gb * inv - b * k; o[2] = gc * inv - c * k;
  }
}

}  // namespace

void launch_activate_forward(int P, const float* scaling, const float* rotation, const float* opacity,
                             const float* feature, float* scales_out, float* rotations_out, float* opacities_out,
                             float* feature_out, cudaStream_t s) {
  if (P <= 0) return;
  activate_forward_kernel<<<(P + ACT_THREADS - 1) / ACT_THREADS, ACT_THREADS, 0, s>>>(
      P, scaling, rotation, opacity, feature, scales_out, rotations_out, opacities_out, feature_out);
  count_launch();
}

void launch_activate_backward(int P, const float* scaling, const float* rotation, const float* opacity,
                              const float* feature, const float* g_scales, const float* g_rotations,
                              const float* g_opacities, const float* g_feature, float* d_scaling, float* d_rotation,
                              float* d_opacity, float* d_feature, cudaStream_t s) {
  if (P <= 0) return;
  activate_backward_kernel<<<(P + ACT_THREADS - 1) / ACT_THREADS, ACT_THREADS, 0, s>>>(
      P, scaling, rotation, opacity, feature, g_scales, g_rotations, g_opacities, g_feature, d_scaling, d_rotation,
      d_opacity, d_feature);
  count_launch();
}

}  // namespace lsr
