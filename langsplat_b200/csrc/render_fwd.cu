// render_fwd.cu — K6': 16x16-tile front-to-back alpha blend of RGB + the 3-channel language feature.
// Replaces FORWARD::render / renderCUDA (RAST/cuda_rasterizer/forward.cu:262-426).
//
// One CTA per tile, one thread per pixel.  Per 256-instance batch the CTA gathers one 48-byte GeomRec per instance
// into shared memory (position, conic, opacity, colour AND feature — the reference re-reads colour/feature from
// global memory for every contributing pair, forward.cu:359-366), then every pixel walks the batch.
// The per-pair arithmetic is written with explicit round-to-nearest intrinsics in the exact order of the reference's
// SASS so that the skip / stop decisions, hence n_contrib, are bit-identical:
//   power = fma(fma(dx, cx*dx, (cz*dy)*dy), -0.5, -((cy*dx)*dy));  alpha = min(0.99, o*expf(power)).
// include_feature is a template parameter (a runtime branch in the reference).
#include "common.cuh"

namespace lsr {

namespace {

template <bool FEAT>
__global__ void __launch_bounds__(TILE_PIX)
render_fwd_kernel(int W, int H, const uint2* __restrict__ ranges, const uint32_t* __restrict__ list,
                  const GeomRec* __restrict__ rec, const float* __restrict__ bg, float* __restrict__ final_T,
                  uint32_t* __restrict__ n_contrib, float* __restrict__ out_color, float* __restrict__ out_feature) {
  __shared__ float4 s_a[TILE_PIX];  // x, y, cx, cy
  __shared__ float4 s_b[TILE_PIX];  // cz, o, r, g
  __shared__ float4 s_c[TILE_PIX];  // b, f0, f1, f2

  const int tx = threadIdx.x & (TILE_X - 1), ty = threadIdx.x >> 4;
  const int tid = threadIdx.x;
  const int tiles_x = (W + TILE_X - 1) / TILE_X;
  const uint32_t px = blockIdx.x * TILE_X + tx, py = blockIdx.y * TILE_Y + ty;
  const bool inside = px < (uint32_t)W && py < (uint32_t)H;
  const float pxf = (float)px, pyf = (float)py;
  const uint2 range = ranges[blockIdx.y * tiles_x + blockIdx.x];
  const int rounds = (int)((range.y - range.x + TILE_PIX - 1) / TILE_PIX);
  int todo = (int)(range.y - range.x);

  bool done = !inside;
  float T = 1.0f;
  uint32_t contributor = 0, last_contributor = 0;
  float C0 = 0.f, C1 = 0.f, C2 = 0.f, F0 = 0.f, F1 = 0.f, F2 = 0.f;

  for (int i = 0; i < rounds; i++, todo -= TILE_PIX) {
    if (__syncthreads_count(done) == TILE_PIX) break;
    const uint32_t progress = (uint32_t)i * TILE_PIX + tid;
    if (range.x + progress < range.y) {
      const uint32_t id = __ldg(list + range.x + progress);
      const float4* r4 = reinterpret_cast<const float4*>(rec + id);
      s_a[tid] = __ldg(r4);
      s_b[tid] = __ldg(r4 + 1);
      s_c[tid] = __ldg(r4 + 2);
    }
    __syncthreads();
    const int nb = min(TILE_PIX, todo);
    for (int j = 0; !done && j < nb; j++) {
      contributor++;
      const float4 a = s_a[j];
      const float4 b = s_b[j];
      const float dx = __fadd_rn(a.x, -pxf), dy = __fadd_rn(a.y, -pyf);
      const float q = __fmaf_rn(dx, __fmul_rn(a.z, dx), __fmul_rn(__fmul_rn(b.x, dy), dy));
      const float power = __fmaf_rn(q, -0.5f, -__fmul_rn(__fmul_rn(a.w, dx), dy));
      if (power > 0.0f) continue;
      const float alpha = fminf(0.99f, __fmul_rn(b.y, expf(power)));
      if (alpha < 1.0f / 255.0f) continue;
      const float test_T = __fmul_rn(T, __fadd_rn(1.0f, -alpha));
      if (test_T < 0.0001f) {
        done = true;
        continue;
      }
      const float4 c = s_c[j];
      C0 = __fmaf_rn(T, __fmul_rn(b.z, alpha), C0);
      C1 = __fmaf_rn(T, __fmul_rn(b.w, alpha), C1);
      C2 = __fmaf_rn(T, __fmul_rn(c.x, alpha), C2);
      if (FEAT) {
        F0 = __fmaf_rn(T, __fmul_rn(c.y, alpha), F0);
        F1 = __fmaf_rn(T, __fmul_rn(c.z, alpha), F1);
        F2 = __fmaf_rn(T, __fmul_rn(c.w, alpha), F2);
      }
      T = test_T;
      last_contributor = contributor;
    }
  }
  if (inside) {
    const size_t N = (size_t)W * H;
    const size_t pix = (size_t)py * W + px;
    final_T[pix] = T;
    n_contrib[pix] = last_contributor;
    out_color[pix] = __fmaf_rn(T, __ldg(bg), C0);
    out_color[N + pix] = __fmaf_rn(T, __ldg(bg + 1), C1);
    out_color[2 * N + pix] = __fmaf_rn(T, __ldg(bg + 2), C2);
    if (FEAT) {
      out_feature[pix] = F0;  // no background term for the feature image (forward.cu:388)
      out_feature[N + pix] = F1;
      out_feature[2 * N + pix] = F2;
    }
  }
}

}  // namespace

void launch_render_fwd(int W, int H, const uint2* ranges, const uint32_t* list, const GeomRec* rec, const float* bg,
                       bool include_feature, float* final_T, uint32_t* n_contrib, float* out_color,
                       float* out_feature, cudaStream_t s) {
  dim3 grid((W + TILE_X - 1) / TILE_X, (H + TILE_Y - 1) / TILE_Y);
  if (grid.x == 0 || grid.y == 0) return;
  if (include_feature)
    render_fwd_kernel<true><<<grid, TILE_PIX, 0, s>>>(W, H, ranges, list, rec, bg, final_T, n_contrib, out_color,
                                                      out_feature);
  else
    render_fwd_kernel<false><<<grid, TILE_PIX, 0, s>>>(W, H, ranges, list, rec, bg, final_T, n_contrib, out_color,
                                                       out_feature);
  count_launch();
}

}  // namespace lsr
