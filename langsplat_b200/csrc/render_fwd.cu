// render_fwd.cu — K6': 16x16-tile front-to-back alpha blend of RGB + the 3-channel language feature.
// Replaces FORWARD::render / renderCUDA (RAST/cuda_rasterizer/forward.cu:262-426).
//
// One single-warp CTA per 8x8-pixel quadrant of a tile, two pixels per thread (rows r and r+4), so the shared-memory
// broadcasts, the x half of the quadratic form and the loop overhead are paid once per 64 pixels, and no CTA barrier
// exists.  Per 32-instance batch the warp gathers one 48-byte GeomRec per instance (position, conic, opacity, colour
// AND feature — the reference re-reads colour/feature from global memory for every contributing pair,
// forward.cu:359-366); instances that provably cannot contribute anywhere in the quadrant are dropped at staging
// (render_common.cuh).  The next batch's records are fetched into registers before the current one is blended.
// The per-pair arithmetic is written with explicit round-to-nearest intrinsics in the exact order of the reference's
// SASS so that the skip / stop decisions, hence n_contrib, are bit-identical:
//   power = fma(fma(dx, cx*dx, (cz*dy)*dy), -0.5, -((cy*dx)*dy));  alpha = min(0.99, o*expf(power)).
// expf is only evaluated when power >= thr_m (see render_common.cuh).  include_feature is a template parameter
// (a runtime branch in the reference).  Kernel is FP32-issue bound; shared-memory reads are warp broadcasts.
#include "render_common.cuh"

namespace lsr {

namespace {

constexpr int FWD_SLOTS = BLEND_SLOTS;

// COLOR = false: feature-only re-blend of a frame whose geometry is already binned (lsr_forward_feature_level): the
// same walk and the same decisions, but no RGB accumulation and no colour / final_T / n_contrib stores.
// WEIGHT: one blend weight alpha * T per pair (opt-in, see fast_blend() below); false = the reference's rounding order
template <bool FEAT, bool COLOR, bool WEIGHT>
__global__ void __launch_bounds__(32)
render_fwd_kernel(int W, int H, int tiles_x, const uint2* __restrict__ ranges, const uint32_t* __restrict__ order,
                  const uint32_t* __restrict__ list,
                  const GeomRec* __restrict__ rec, const float* __restrict__ bg, float* __restrict__ final_T,
                  uint32_t* __restrict__ n_contrib, float* __restrict__ out_color, float* __restrict__ out_feature) {
#if LSR_FWD_PK
  __shared__ WarpBatchPk sm;
#else
  __shared__ WarpBatch sm;
#endif

  const int lane = threadIdx.x;
  const uint32_t quad = blockIdx.x % QUADS;
  const uint32_t tile = order ? __ldg(order + blockIdx.x / QUADS) : blockIdx.x / QUADS;
  const uint32_t qx0 = (tile % tiles_x) * TILE_X + (quad % QUADS_X) * QUAD_X;
  const uint32_t qy0 = (tile / tiles_x) * TILE_Y + (quad / QUADS_X) * QUAD_Y;
  const uint32_t px = qx0 + (lane & 7), py0 = qy0 + (lane >> 3);  // slot s adds 4*s rows
  const float pxf = (float)px;
  const float X0 = (float)qx0, Y0 = (float)qy0;
  const uint2 range = ranges[tile];
  const int len = (int)(range.y - range.x);

  // lim[s] is the bound of the reference's `power > 0` skip (forward.cu:343); once pixel s has finished it drops to
  // -inf, which retires the pixel inside the same compare (a NaN power still gets through and is caught below).
  float pyf[FWD_SLOTS], T[FWD_SLOTS], lim[FWD_SLOTS], C[FWD_SLOTS][3], F[FWD_SLOTS][3];
  bool inside[FWD_SLOTS];
  uint32_t last_contributor[FWD_SLOTS];
  const float kDone = __int_as_float(0xff800000);
#pragma unroll
  for (int s = 0; s < FWD_SLOTS; s++) {
    const uint32_t py = py0 + 4 * s;
    pyf[s] = (float)py;
    inside[s] = px < (uint32_t)W && py < (uint32_t)H;
    lim[s] = inside[s] ? 0.0f : kDone;
    T[s] = 1.0f;
    last_contributor[s] = 0;
#pragma unroll
    for (int ch = 0; ch < 3; ch++) { C[s][ch] = 0.f; F[s][ch] = 0.f; }
  }

  const uint32_t smA = smem_u32(&sm.A[0]);
#if LSR_FWD_PK
  const f32x2 npx2 = pk2(-pxf, -pxf), npy2 = pk2(-pyf[0], -pyf[1]);
#endif
  Fetched nxt = fetch_instance(list, rec, range.x, lane < len ? lane : -1);
  for (int base = 0; base < len; base += 32) {
    // stop when no pixel of the quadrant is still blending; also ends everyone's reads of the previous batch
    if (__all_sync(0xffffffffu, lim[0] < 0.0f && lim[1] < 0.0f)) break;
    const int nb = commit_batch(sm, nxt, X0, Y0);
    {
      const int li = base + 32 + lane;
      nxt = fetch_instance(list, rec, range.x, li < len ? li : -1);
    }
    for (int j = 0; j < nb; j++) {
      const uint32_t ad = smA + (uint32_t)j * (uint32_t)sizeof(float4);
      float power[FWD_SLOTS];
      bool ok[FWD_SLOTS];
#if LSR_FWD_PK
      // both pixels of the thread at once; every packed op is the scalar path's op on each half (same roundings)
      const float4 a = lds128(ad);            // x, x, y, y
      const float4 q4 = lds128(ad + SM_B);    // cx, cx, -cy, -cy
      const float4 b4 = lds128(ad + 2 * SM_B);  // cz, cz, thr_m, o
      const f32x2 dxp = add2(pk2(a.x, a.y), npx2);
      const f32x2 dyp = add2(pk2(a.z, a.w), npy2);
      const f32x2 qxp = mul2(pk2(q4.x, q4.y), dxp);
      const f32x2 nhxp = mul2(pk2(q4.z, q4.w), dxp);  // -(cy*dx)
      const f32x2 qq = fma2(dxp, qxp, mul2(mul2(pk2(b4.x, b4.y), dyp), dyp));
      const f32x2 pw = fma2(qq, pk2(-0.5f, -0.5f), mul2(nhxp, dyp));
      upk2(pw, power[0], power[1]);
      const float thr_m = b4.z, opac = b4.w;
#pragma unroll
      for (int s = 0; s < FWD_SLOTS; s++) ok[s] = !(power[s] > lim[s]) && !(power[s] < thr_m);
      if (!(ok[0] || ok[1])) continue;
      const float4 c = lds128(ad + 3 * SM_B);
      const float4 d = lds128(ad + 4 * SM_B);  // f1, f2, pos, id
      const float pos_bits = d.z;
#else
      const float4 a = lds128(ad);
      const float4 b = lds128(ad + SM_B);
      const float dx = __fadd_rn(a.x, -pxf);
      const float qx = __fmul_rn(a.z, dx), hx = __fmul_rn(a.w, dx);
#pragma unroll
      for (int s = 0; s < FWD_SLOTS; s++) {
        const float dy = __fadd_rn(a.y, -pyf[s]);
        const float q = __fmaf_rn(dx, qx, __fmul_rn(__fmul_rn(b.x, dy), dy));
        power[s] = __fmaf_rn(q, -0.5f, -__fmul_rn(hx, dy));
        // b.y = thr_m: below it alpha < 1/255 for certain (a NaN power blends, as in the reference)
        ok[s] = !(power[s] > lim[s]) && !(power[s] < b.y);
      }
      if (!(ok[0] || ok[1])) continue;
      const float opac = b.z, pos_bits = b.w;
      const float4 c = lds128(ad + SM_C);
      float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
      if (FEAT) d = lds128(ad + SM_D);
#endif
#pragma unroll
      for (int s = 0; s < FWD_SLOTS; s++) {
        if (!ok[s] || lim[s] < 0.0f) continue;
        const float alpha = fminf(0.99f, __fmul_rn(opac, expf(power[s])));
        if (alpha < kAlphaMin) continue;
        const float test_T = __fmul_rn(T[s], __fadd_rn(1.0f, -alpha));
        if (test_T < 0.0001f) {
          lim[s] = kDone;  // done (forward.cu:353-357)
          continue;
        }
        if (WEIGHT) {
          // One blend weight per pair, w = alpha * T, instead of (channel * alpha) * T per channel
          // (forward.cu:361-366): 7 instead of 12 instructions for the six channels.  Each term then differs from the
          // reference's by at most one rounding; alpha, T and every decision are untouched.
          const float w = __fmul_rn(alpha, T[s]);
          if (COLOR) {
            C[s][0] = __fmaf_rn(w, c.x, C[s][0]);
            C[s][1] = __fmaf_rn(w, c.y, C[s][1]);
            C[s][2] = __fmaf_rn(w, c.z, C[s][2]);
          }
          if (FEAT) {
            F[s][0] = __fmaf_rn(w, c.w, F[s][0]);
            F[s][1] = __fmaf_rn(w, d.x, F[s][1]);
            F[s][2] = __fmaf_rn(w, d.y, F[s][2]);
          }
        } else {
          if (COLOR) {
            C[s][0] = __fmaf_rn(T[s], __fmul_rn(c.x, alpha), C[s][0]);
            C[s][1] = __fmaf_rn(T[s], __fmul_rn(c.y, alpha), C[s][1]);
            C[s][2] = __fmaf_rn(T[s], __fmul_rn(c.z, alpha), C[s][2]);
          }
          if (FEAT) {
            F[s][0] = __fmaf_rn(T[s], __fmul_rn(c.w, alpha), F[s][0]);
            F[s][1] = __fmaf_rn(T[s], __fmul_rn(d.x, alpha), F[s][1]);
            F[s][2] = __fmaf_rn(T[s], __fmul_rn(d.y, alpha), F[s][2]);
          }
        }
        T[s] = test_T;
        last_contributor[s] = __float_as_uint(pos_bits) + 1u;  // 1-based position in the tile's list (forward.cu:372)
      }
    }
  }
  const size_t N = (size_t)W * H;
#pragma unroll
  for (int s = 0; s < FWD_SLOTS; s++) {
    if (!inside[s]) continue;
    const size_t pix = (size_t)(py0 + 4 * s) * W + px;
    if (COLOR) {
      final_T[pix] = T[s];
      n_contrib[pix] = last_contributor[s];
      out_color[pix] = __fmaf_rn(T[s], __ldg(bg), C[s][0]);
      out_color[N + pix] = __fmaf_rn(T[s], __ldg(bg + 1), C[s][1]);
      out_color[2 * N + pix] = __fmaf_rn(T[s], __ldg(bg + 2), C[s][2]);
    }
    if (FEAT) {
      out_feature[pix] = F[s][0];  // no background term for the feature image (forward.cu:388)
      out_feature[N + pix] = F[s][1];
      out_feature[2 * N + pix] = F[s][2];
    }
  }
}


// language feature of every visible Gaussian's record replaced by `lang` [P,3]
__global__ void __launch_bounds__(256)
refeature_kernel(int P, const int32_t* __restrict__ radii, const float* __restrict__ lang, GeomRec* __restrict__ rec) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P || radii[idx] <= 0) return;
  rec[idx].f0 = __ldg(lang + 3 * (size_t)idx);
  rec[idx].f1 = __ldg(lang + 3 * (size_t)idx + 1);
  rec[idx].f2 = __ldg(lang + 3 * (size_t)idx + 2);
}

}  // namespace

// With the reference's rounding order the images are BIT-IDENTICAL to the reference extension's (measured at C2 and
// C3, profiles/r2ae/image_diff_main.log: 0 of 6.2 M pixels differ).  LSR_FAST_BLEND=1 (or a build with
// -DLSR_FWD_WEIGHT=1) trades that for speed: render_fwd 440 -> 408 us at C2, 601.7 -> 612.1 frames/s, C5 458 -> 470;
// a quarter of the pixels then differ by one or two ulp (max 2.4e-7, the images' bar is 1e-5).
static bool fast_blend() {
  static const bool on = LSR_FWD_WEIGHT != 0 || (getenv("LSR_FAST_BLEND") != nullptr && getenv("LSR_FAST_BLEND")[0] == '1');
  return on;
}

void launch_render_fwd(int W, int H, const uint2* ranges, const uint32_t* order, const uint32_t* list, const GeomRec* rec, const float* bg,
                       bool include_feature, float* final_T, uint32_t* n_contrib, float* out_color,
                       float* out_feature, cudaStream_t s) {
  const int tiles_x = (W + TILE_X - 1) / TILE_X, tiles_y = (H + TILE_Y - 1) / TILE_Y;
  if (tiles_x == 0 || tiles_y == 0) return;
  const unsigned grid = (unsigned)tiles_x * (unsigned)tiles_y * QUADS;
#define LSR_FWD(F, C, WT) render_fwd_kernel<F, C, WT><<<grid, 32, 0, s>>>(W, H, tiles_x, ranges, order, list, rec, bg, \
                                                                        final_T, n_contrib, out_color, out_feature)
  if (include_feature) { if (fast_blend()) LSR_FWD(true, true, true); else LSR_FWD(true, true, false); }
  else                 { if (fast_blend()) LSR_FWD(false, true, true); else LSR_FWD(false, true, false); }
#undef LSR_FWD
  count_launch();
}

void launch_refeature_render(int P, int W, int H, const int32_t* radii, const float* lang, GeomRec* rec,
                             const uint2* ranges, const uint32_t* order, const uint32_t* list, float* out_feature,
                             cudaStream_t s) {
  const int tiles_x = (W + TILE_X - 1) / TILE_X, tiles_y = (H + TILE_Y - 1) / TILE_Y;
  if (P > 0) {
    refeature_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, radii, lang, rec);
    count_launch();
  }
  if (tiles_x == 0 || tiles_y == 0) return;
  const unsigned grid = (unsigned)tiles_x * (unsigned)tiles_y * QUADS;
  if (fast_blend())
    render_fwd_kernel<true, false, true><<<grid, 32, 0, s>>>(W, H, tiles_x, ranges, order, list, rec, nullptr, nullptr,
                                                             nullptr, nullptr, out_feature);
  else
    render_fwd_kernel<true, false, false><<<grid, 32, 0, s>>>(W, H, tiles_x, ranges, order, list, rec, nullptr, nullptr,
                                                              nullptr, nullptr, out_feature);
  count_launch();
}

}  // namespace lsr
