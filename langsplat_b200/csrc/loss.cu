// loss.cu — fused (masked) L1 loss on the rendered images: the step right after the rasterizer in LangSplat's
// training loop.  Replaces, for pred/gt [C,N] and an optional mask [N] broadcast over C,
//     l1_loss(pred * mask, gt * mask) = torch.abs(pred * mask - gt * mask).mean()
// (utils/loss_utils.py:17-18 as called from train.py:96-99), which PyTorch runs as five elementwise / reduction
// kernels forward and four backward, each streaming a [C,H,W] temporary through HBM.  Here the forward is one pass
// (reads pred, gt, mask; one pre-scaled atomicAdd per CTA) and the backward one pass (reads pred, gt, mask and the
// upstream scalar; writes dL_dpred = g * sign(pred*m - gt*m) * m / (C*N)).  HBM-bound: 8 B/element forward,
// 12 B/element backward (+ 4/C B for the mask).
// Arithmetic follows PyTorch's: the two products are rounded separately before the subtraction.
// The ground truth may also be given as the 8-bit image it was stored as (GT = uint8_t): the kernels then form
// gt = u / 255 themselves (IEEE division, the value utils/general_utils.py:23 PILtoTorch computes on the host), so the
// float copy of the image is never written to HBM: 5 B/element forward, 9 B/element backward.
#include "common.cuh"

namespace lsr {

namespace {

constexpr int LOSS_THREADS = 256;
constexpr int LOSS_VEC_PER_THREAD = 4;  // float4 per thread

__device__ __forceinline__ float diff_masked(float p, float g, float m) {
  return __fsub_rn(__fmul_rn(p, m), __fmul_rn(g, m));
}

__device__ __forceinline__ float gt1(const float* gt, int64_t i) { return __ldg(gt + i); }
__device__ __forceinline__ float gt1(const uint8_t* gt, int64_t i) { return __fdiv_rn((float)__ldg(gt + i), 255.0f); }
__device__ __forceinline__ float4 gt4(const float* gt, int64_t v) { return __ldg(reinterpret_cast<const float4*>(gt) + v); }
__device__ __forceinline__ float4 gt4(const uint8_t* gt, int64_t v) {
  const uchar4 u = __ldg(reinterpret_cast<const uchar4*>(gt) + v);
  return make_float4(__fdiv_rn((float)u.x, 255.0f), __fdiv_rn((float)u.y, 255.0f), __fdiv_rn((float)u.z, 255.0f),
                     __fdiv_rn((float)u.w, 255.0f));
}

template <bool MASK, typename GT>
__global__ void __launch_bounds__(LOSS_THREADS)
l1_forward_kernel(const float* __restrict__ pred, const GT* __restrict__ gt, const float* __restrict__ mask,
                  int64_t N, int64_t total, float inv_count, bool vec_ok, float* __restrict__ loss) {
  __shared__ float s_part[LOSS_THREADS / 32];
  const int64_t nvec = total / 4;
  float acc = 0.f;
  // vec_ok (host): N % 4 == 0 — a float4 never straddles two channels, so the mask index is (i % N) — and every
  // pointer is 16-byte aligned (a contiguous view may start at any element of its storage)
  if (vec_ok) {
    const float4* p4 = reinterpret_cast<const float4*>(pred);
    const int64_t nv_per_ch = N / 4;
    for (int64_t v = (int64_t)blockIdx.x * LOSS_THREADS * LOSS_VEC_PER_THREAD + threadIdx.x, it = 0;
         it < LOSS_VEC_PER_THREAD; it++, v += LOSS_THREADS) {
      if (v >= nvec) break;
      const float4 p = __ldg(p4 + v), g = gt4(gt, v);
      float4 m = make_float4(1.f, 1.f, 1.f, 1.f);
      if (MASK) m = __ldg(reinterpret_cast<const float4*>(mask) + (v % nv_per_ch));
      acc += fabsf(diff_masked(p.x, g.x, m.x)) + fabsf(diff_masked(p.y, g.y, m.y)) +
             fabsf(diff_masked(p.z, g.z, m.z)) + fabsf(diff_masked(p.w, g.w, m.w));
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * LOSS_THREADS * LOSS_VEC_PER_THREAD * 4 + threadIdx.x, it = 0;
         it < LOSS_VEC_PER_THREAD * 4; it++, i += LOSS_THREADS) {
      if (i >= total) break;
      const float m = MASK ? __ldg(mask + (i % N)) : 1.f;
      acc += fabsf(diff_masked(__ldg(pred + i), gt1(gt, i), m));
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < LOSS_THREADS / 32; w++) t += s_part[w];
    atomicAdd(loss, t * inv_count);
  }
}

template <bool MASK, typename GT>
__global__ void __launch_bounds__(LOSS_THREADS)
l1_backward_kernel(const float* __restrict__ pred, const GT* __restrict__ gt, const float* __restrict__ mask,
                   int64_t N, int64_t total, float inv_count, bool vec_ok, const float* __restrict__ grad_out,
                   float* __restrict__ dL_dpred) {
  const float gscale = __ldg(grad_out) * inv_count;
  // d|x|/dx = sign(x) with sign(0) = 0, as torch.abs' backward
  auto grad1 = [&](float p, float g, float m) {
    const float d = diff_masked(p, g, m);
    const float sg = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
    return gscale * sg * m;
  };
  const int64_t nvec = total / 4;
  if (vec_ok) {
    const float4* p4 = reinterpret_cast<const float4*>(pred);
    float4* o4 = reinterpret_cast<float4*>(dL_dpred);
    const int64_t nv_per_ch = N / 4;
    for (int64_t v = (int64_t)blockIdx.x * LOSS_THREADS * LOSS_VEC_PER_THREAD + threadIdx.x, it = 0;
         it < LOSS_VEC_PER_THREAD; it++, v += LOSS_THREADS) {
      if (v >= nvec) break;
      const float4 p = __ldg(p4 + v), g = gt4(gt, v);
      float4 m = make_float4(1.f, 1.f, 1.f, 1.f);
      if (MASK) m = __ldg(reinterpret_cast<const float4*>(mask) + (v % nv_per_ch));
      o4[v] = make_float4(grad1(p.x, g.x, m.x), grad1(p.y, g.y, m.y), grad1(p.z, g.z, m.z), grad1(p.w, g.w, m.w));
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * LOSS_THREADS * LOSS_VEC_PER_THREAD * 4 + threadIdx.x, it = 0;
         it < LOSS_VEC_PER_THREAD * 4; it++, i += LOSS_THREADS) {
      if (i >= total) break;
      const float m = MASK ? __ldg(mask + (i % N)) : 1.f;
      dL_dpred[i] = grad1(__ldg(pred + i), gt1(gt, i), m);
    }
  }
}

inline bool aligned16(const void* a, const void* b, const void* c, const void* d = nullptr) {
  return ((((uintptr_t)a) | ((uintptr_t)b) | ((uintptr_t)c) | ((uintptr_t)d)) & 15) == 0;
}

inline unsigned loss_blocks(int64_t total) {
  const int64_t per = (int64_t)LOSS_THREADS * LOSS_VEC_PER_THREAD * 4;
  return (unsigned)((total + per - 1) / per);
}

template <typename GT>
cudaError_t l1_forward(const float* pred, const GT* gt, const float* mask, int C, int64_t N, float* loss,
                       cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(loss, 0, sizeof(float), s);
  if (e != cudaSuccess) return e;
  const int64_t total = (int64_t)C * N;
  if (total == 0) return cudaSuccess;
  const float inv = (float)(1.0 / (double)total);
  // (uchar4 loads of an 8-bit ground truth need 4-byte alignment only)
  const bool vec = (N % 4) == 0 && aligned16(pred, mask, nullptr) && (((uintptr_t)gt) & (sizeof(GT) * 4 - 1)) == 0;
  if (mask) l1_forward_kernel<true, GT><<<loss_blocks(total), LOSS_THREADS, 0, s>>>(pred, gt, mask, N, total, inv, vec, loss);
  else l1_forward_kernel<false, GT><<<loss_blocks(total), LOSS_THREADS, 0, s>>>(pred, gt, mask, N, total, inv, vec, loss);
  count_launch();
  return cudaGetLastError();
}

template <typename GT>
cudaError_t l1_backward(const float* pred, const GT* gt, const float* mask, int C, int64_t N, const float* grad_out,
                        float* dL_dpred, cudaStream_t s) {
  const int64_t total = (int64_t)C * N;
  if (total == 0) return cudaSuccess;
  const float inv = (float)(1.0 / (double)total);
  const bool vec = (N % 4) == 0 && aligned16(pred, mask, dL_dpred) && (((uintptr_t)gt) & (sizeof(GT) * 4 - 1)) == 0;
  if (mask)
    l1_backward_kernel<true, GT><<<loss_blocks(total), LOSS_THREADS, 0, s>>>(pred, gt, mask, N, total, inv, vec,
                                                                             grad_out, dL_dpred);
  else
    l1_backward_kernel<false, GT><<<loss_blocks(total), LOSS_THREADS, 0, s>>>(pred, gt, mask, N, total, inv, vec,
                                                                              grad_out, dL_dpred);
  count_launch();
  return cudaGetLastError();
}

}  // namespace

cudaError_t launch_l1_forward(const float* pred, const float* gt, const float* mask, int C, int64_t N, float* loss,
                              cudaStream_t s) {
  return l1_forward(pred, gt, mask, C, N, loss, s);
}
cudaError_t launch_l1_forward_u8(const float* pred, const uint8_t* gt, const float* mask, int C, int64_t N, float* loss,
                                 cudaStream_t s) {
  return l1_forward(pred, gt, mask, C, N, loss, s);
}
cudaError_t launch_l1_backward(const float* pred, const float* gt, const float* mask, int C, int64_t N,
                               const float* grad_out, float* dL_dpred, cudaStream_t s) {
  return l1_backward(pred, gt, mask, C, N, grad_out, dL_dpred, s);
}
cudaError_t launch_l1_backward_u8(const float* pred, const uint8_t* gt, const float* mask, int C, int64_t N,
                                  const float* grad_out, float* dL_dpred, cudaStream_t s) {
  return l1_backward(pred, gt, mask, C, N, grad_out, dL_dpred, s);
}

}  // namespace lsr
