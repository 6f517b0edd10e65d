// render_fwd.cu — K6': 16x16-tile front-to-back alpha blend of RGB + the 3-channel language feature.
// Replaces FORWARD::render / renderCUDA (RAST/cuda_rasterizer/forward.cu:262-426).
//
// One CTA per tile, one thread per pixel.  Per 256-instance batch the CTA gathers one 48-byte GeomRec per instance
// (position, conic, opacity, colour AND feature — the reference re-reads colour/feature from global memory for every
// contributing pair, forward.cu:359-366), drops instances that provably cannot contribute anywhere in the tile and
// compacts the rest into shared memory (render_common.cuh); then every pixel walks the survivors.
// The per-pair arithmetic is written with explicit round-to-nearest intrinsics in the exact order of the reference's
// SASS so that the skip / stop decisions, hence n_contrib, are bit-identical:
//   power = fma(fma(dx, cx*dx, (cz*dy)*dy), -0.5, -((cy*dx)*dy));  alpha = min(0.99, o*expf(power)).
// expf is only evaluated when power >= thr_m (see render_common.cuh).  include_feature is a template parameter
// (a runtime branch in the reference).  Kernel is FP32-issue bound; shared-memory reads are warp broadcasts.
#include "render_common.cuh"

namespace lsr {

namespace {

template <bool FEAT>
__global__ void __launch_bounds__(TILE_PIX)
render_fwd_kernel(int W, int H, const uint2* __restrict__ ranges, const uint32_t* __restrict__ list,
                  const GeomRec* __restrict__ rec, const float* __restrict__ bg, float* __restrict__ final_T,
                  uint32_t* __restrict__ n_contrib, float* __restrict__ out_color, float* __restrict__ out_feature) {
  __shared__ BatchSmem sm;

  const int tid = threadIdx.x;
  // warp w owns an 8x4-pixel footprint (compact, instead of the reference's 16x2 strips)
  const int tx = ((tid >> 5) & 1) * 8 + (tid & 7), ty = (tid >> 6) * 4 + ((tid & 31) >> 3);
  const int tiles_x = (W + TILE_X - 1) / TILE_X;
  const uint32_t px = blockIdx.x * TILE_X + tx, py = blockIdx.y * TILE_Y + ty;
  const bool inside = px < (uint32_t)W && py < (uint32_t)H;
  const float pxf = (float)px, pyf = (float)py;
  const float X0 = (float)(blockIdx.x * TILE_X), Y0 = (float)(blockIdx.y * TILE_Y);
  const uint2 range = ranges[blockIdx.y * tiles_x + blockIdx.x];
  const int len = (int)(range.y - range.x);
  const int rounds = (len + TILE_PIX - 1) / TILE_PIX;

  bool done = !inside;
  float T = 1.0f;
  uint32_t last_contributor = 0;
  float C0 = 0.f, C1 = 0.f, C2 = 0.f, F0 = 0.f, F1 = 0.f, F2 = 0.f;

  for (int i = 0; i < rounds; i++) {
    if (__syncthreads_count(done) == TILE_PIX) break;
    const int li = i * TILE_PIX + tid;
    const int nb = stage_batch<TILE_PIX>(sm, list, rec, range.x, li < len ? li : -1, X0, Y0);
    for (int j = 0; !done && j < nb; j++) {
      const float4 a = sm.A[j];
      const float4 b = sm.B[j];
      const float dx = __fadd_rn(a.x, -pxf), dy = __fadd_rn(a.y, -pyf);
      const float q = __fmaf_rn(dx, __fmul_rn(a.z, dx), __fmul_rn(__fmul_rn(b.x, dy), dy));
      const float power = __fmaf_rn(q, -0.5f, -__fmul_rn(__fmul_rn(a.w, dx), dy));
      if (power > 0.0f || power < b.y) continue;  // b.y = thr_m: below it alpha < 1/255 for certain
      const float alpha = fminf(0.99f, __fmul_rn(b.z, expf(power)));
      if (alpha < kAlphaMin) continue;
      const float test_T = __fmul_rn(T, __fadd_rn(1.0f, -alpha));
      if (test_T < 0.0001f) {
        done = true;
        continue;
      }
      const float4 c = sm.C[j];
      C0 = __fmaf_rn(T, __fmul_rn(c.x, alpha), C0);
      C1 = __fmaf_rn(T, __fmul_rn(c.y, alpha), C1);
      C2 = __fmaf_rn(T, __fmul_rn(c.z, alpha), C2);
      if (FEAT) {
        const float4 d = sm.D[j];
        F0 = __fmaf_rn(T, __fmul_rn(c.w, alpha), F0);
        F1 = __fmaf_rn(T, __fmul_rn(d.x, alpha), F1);
        F2 = __fmaf_rn(T, __fmul_rn(d.y, alpha), F2);
      }
      T = test_T;
      last_contributor = __float_as_uint(b.w) + 1u;  // 1-based position in the tile's list (forward.cu:372)
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
