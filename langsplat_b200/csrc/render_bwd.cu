// render_bwd.cu — K7': per-pixel reverse blend; gradients w.r.t. mean2D, conic, opacity, colour, language feature.
// Replaces BACKWARD::render / renderCUDA (RAST/cuda_rasterizer/backward.cu:399-593).
//
// The kernel is instruction-issue bound (ncu: 85 % issue-active), so everything here is about instructions per
// contributing (pixel, Gaussian) pair:
//  * Reduction instead of atomics.  The reference issues 12 global atomicAdds per contributing pair
//    (backward.cu:541-590).  Here all lanes of a warp walk the same instance in lock-step, so the (up to) 12 per-lane
//    contributions are summed across the warp with a value-halving butterfly (13 shuffles for 12 values instead of
//    60: each round a lane keeps half of its values and sends the other half to its partner), leaving each value's
//    warp total on one lane, which issues a single RED.ADD.F32.
//  * Two pixels per thread.  A warp owns an 8x8 pixel quadrant (and is a CTA of its own: no barriers); each thread
//    blends rows r and r+4 of its column and adds both pixels' contributions before the butterfly, so the reduction,
//    the shared-memory broadcasts and the loop overhead are paid once per 64 pixels (measured on the bench scene: 3.6 M
//    working (warp, instance) pairs instead of 6.2 M with 8x4 footprints).
//  * Scalar "colour behind" recurrence.  The reference keeps accum_rec[ch] / last_color[ch] per channel
//    (backward.cu:527-538) only to form sum_ch (c[ch] - accum_rec[ch]) * dL_dpixel[ch]; the recurrence is linear, so
//    it is carried for the dot product itself: A <- last_alpha * (last_color . dL_dpixel) + (1 - last_alpha) * A.
//    One scalar recurrence serves the three colour and the three feature channels (6 FMAs + 3 instead of ~30).
//  * T is walked back with a reciprocal (MUFU.RCP) instead of the reference's IEEE division; gradients carry a 1e-4
//    relative tolerance (the reference's own sums are unordered atomics), the forward decisions are untouched.
//  * The walk starts at the quadrant's maximum of n_contrib, so the never-blended tail of the tile's list is not
//    fetched; instances that cannot reach alpha >= 1/255 anywhere in the quadrant are dropped at staging
//    (render_common.cuh); the next batch's records are fetched while the current one is processed.
//  * Moments instead of gradients (GEOM).  With s = G * dL_dalpha per pair, the reference's six geometric sums
//    (backward.cu:552-582) are linear in the moments  M0 = sum s, Mx = sum s*dx, My = sum s*dy, Mxx = sum s*dx*dx,
//    Mxy = sum s*dx*dy, Myy = sum s*dy*dy  of the pixel offsets about the Gaussian centre, with per-Gaussian
//    coefficients (opacity, conic).  The kernel accumulates the moments (9 instructions per pixel instead of 21) into
//    the dL_dopacity / dL_dmean2D / dL_dconic buffers, and the per-Gaussian kernel that follows (preprocess_bwd.cu)
//    applies the coefficients once per Gaussian:  dL_dopacity = M0,  dL_dmean2D = -o * (cx*Mx + cy*My) * W/2, ...,
//    dL_dconic = -0.5 * o * (Mxx, Mxy, Myy).
// Template switches specialise on what the caller needs (ctx.needs_input_grad): FEAT = language feature blended,
// DCOL = a colour upstream gradient exists, GEOM = conic/opacity gradients wanted.  LangSplat's feature-only training
// (train.py:96-99 with gaussian_model.py:212-217) runs <FEAT=1, DCOL=0, GEOM=0>: 5 values instead of 12.
#include <cstdlib>

#include "render_common.cuh"

namespace lsr {

namespace {

constexpr int SPARSE_LANES = 2;

// Butterfly reduction of N per-lane values; afterwards lane `l` with held_index<N>(l) >= 0 holds in v[0] the warp
// total of value held_index(l).
template <int N, int OFF>
__device__ __forceinline__ void multi_reduce(float (&v)[12], int lane) {
  if constexpr (OFF >= 1) {
    if constexpr (N == 1) {
      v[0] += __shfl_xor_sync(0xffffffffu, v[0], OFF);
      multi_reduce<1, OFF / 2>(v, lane);
    } else {
      constexpr int Hh = (N + 1) / 2;
      const bool upper = (lane & OFF) != 0;
#pragma unroll
      for (int k = 0; k < Hh; k++) {
        const float a = v[k];
        const float b = (k + Hh < N) ? v[k + Hh] : 0.f;
        const float send = upper ? a : b;
        const float keep = upper ? b : a;
        v[k] = keep + __shfl_xor_sync(0xffffffffu, send, OFF);
      }
      multi_reduce<Hh, OFF / 2>(v, lane);
    }
  }
}
template <int N, int OFF>
__device__ __forceinline__ int held_index(int lane) {
  if constexpr (OFF < 1) {
    return 0;
  } else if constexpr (N == 1) {
    // plain rounds: every lane of the pair ends with the total; elect the lower one
    const int rest = held_index<1, OFF / 2>(lane);
    return (lane & OFF) ? -1 : rest;
  } else {
    constexpr int Hh = (N + 1) / 2;
    const int rest = held_index<Hh, OFF / 2>(lane);
    if (rest < 0) return -1;
    const int p = rest + ((lane & OFF) ? Hh : 0);
    return p < N ? p : -1;
  }
}

constexpr int BWD_SLOTS = BLEND_SLOTS;  // pixels per thread (rows r, r+4)

#ifndef LSR_BWD_MINB
#define LSR_BWD_MINB 25  // caps ptxas at 80 registers: 25 instead of 23 warps per SM (measured: 958 -> 905 us)
#endif
template <bool FEAT, bool DCOL, bool GEOM>
__global__ void __launch_bounds__(32, LSR_BWD_MINB)
render_bwd_kernel(int W, int H, int tiles_x, const uint2* __restrict__ ranges, const uint32_t* __restrict__ order,
                  const uint32_t* __restrict__ list,
                  const GeomRec* __restrict__ rec, const float* __restrict__ bg, const float* __restrict__ final_Ts,
                  const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
                  const float* __restrict__ dL_dpixels_F, float* __restrict__ dL_dmean2D,
                  float* __restrict__ dL_dconic, float* __restrict__ dL_dopacity, float* __restrict__ dL_dcolors,
                  float* __restrict__ dL_dfeat) {
  constexpr int I_MEAN = 0;
  constexpr int I_CONIC = 2;
  constexpr int I_OPAC = 5;
  constexpr int I_COL = GEOM ? 6 : 2;
  constexpr int I_FEAT = I_COL + (DCOL ? 3 : 0);
  constexpr int NV = I_FEAT + (FEAT ? 3 : 0);

  // warp reduction through shared memory for wide value sets: lane l stores its NV partials as column l of a
  // [NV][36]-float array (rows padded to 36 so the 128-bit row reads below are conflict-free), then lane (k, h) =
  // (l & 15, l >> 4) sums the 16 partials of half h of row k, one shuffle joins the halves and lane k issues the RED.
  // 12 STS + 4 LDS.128 + 16 FADD + 1 SHFL instead of the butterfly's 13 SHFL + 24 FSEL + 13 FADD.
  constexpr bool SMEM_REDUCE = NV > 6;
  constexpr int RED_ROW = 36;
  __shared__ WarpBatch sm;
  __shared__ __align__(16) float s_red[SMEM_REDUCE ? NV * RED_ROW : 4];

  const int lane = threadIdx.x;
  const uint32_t quad = blockIdx.x % QUADS;
  const uint32_t tile = order ? __ldg(order + blockIdx.x / QUADS) : blockIdx.x / QUADS;
  const uint32_t qx0 = (tile % tiles_x) * TILE_X + (quad % QUADS_X) * QUAD_X;
  const uint32_t qy0 = (tile / tiles_x) * TILE_Y + (quad / QUADS_X) * QUAD_Y;
  const uint32_t px = qx0 + (lane & 7), py0 = qy0 + (lane >> 3);  // slot s adds 4*s rows
  const float pxf = (float)px;
  const float X0 = (float)qx0, Y0 = (float)qy0;
  const size_t N = (size_t)W * H;
  const uint2 range = ranges[tile];

  float pyf[BWD_SLOTS], T[BWD_SLOTS], bgT[BWD_SLOTS];
  uint32_t lastc[BWD_SLOTS];
  float dpix[BWD_SLOTS][3], dpixF[BWD_SLOTS][3];
  float accA[BWD_SLOTS], cd_last[BWD_SLOTS], alpha_last[BWD_SLOTS];
  uint32_t wmax = 0;
#pragma unroll
  for (int s = 0; s < BWD_SLOTS; s++) {
    const uint32_t py = py0 + 4 * s;
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const size_t pix = (size_t)py * W + px;
    pyf[s] = (float)py;
    const float T_final = inside ? final_Ts[pix] : 0.f;
    T[s] = T_final;
    lastc[s] = inside ? n_contrib[pix] : 0u;
    wmax = max(wmax, lastc[s]);
#pragma unroll
    for (int i = 0; i < 3; i++) {
      dpix[s][i] = (DCOL && inside) ? dL_dpixels[i * N + pix] : 0.f;
      dpixF[s][i] = (FEAT && inside) ? dL_dpixels_F[i * N + pix] : 0.f;
    }
    bgT[s] = 0.f;
    if (DCOL) bgT[s] = -T_final * (__ldg(bg) * dpix[s][0] + __ldg(bg + 1) * dpix[s][1] + __ldg(bg + 2) * dpix[s][2]);
    accA[s] = 0.f; cd_last[s] = 0.f; alpha_last[s] = 0.f;
  }
  // The quadrant's maximum of n_contrib bounds the part of the list anyone here needs.
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, d));
  const int eff_len = (int)wmax;  // forward indices [0, eff_len) are the only ones examined

  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;

  // which reduced value this lane ends up holding, and where it goes
  const int held = SMEM_REDUCE ? (((lane & 15) < NV && lane < 16) ? (lane & 15) : -1) : held_index<NV, 16>(lane);
  float* out_base = nullptr;
  int out_stride = 0;
  if (held >= 0) {
    if (held < I_MEAN + 2) { out_base = dL_dmean2D + held; out_stride = 3; }
    else if (GEOM && held < I_OPAC) { const int k = held - I_CONIC; out_base = dL_dconic + (k == 2 ? 3 : k); out_stride = 4; }
    else if (GEOM && held == I_OPAC) { out_base = dL_dopacity; out_stride = 1; }
    else if (DCOL && held >= I_COL && held < I_COL + 3) { out_base = dL_dcolors + (held - I_COL); out_stride = 3; }
    else if (FEAT && held >= I_FEAT) { out_base = dL_dfeat + (held - I_FEAT); out_stride = 3; }
  }

  // back to front: the batch starting at `top` covers forward indices top-lane, top = eff_len-1 downwards
  // (backward.cu:486-498)
  const uint32_t smA = smem_u32(&sm.A[0]);
  // reduction scratch: lane l stores into column l, then reads 16 floats of row min(l & 15, NV-1), half l >> 4
  const uint32_t red_st = smem_u32(s_red) + (uint32_t)lane * 4u;
  const uint32_t red_ld = smem_u32(s_red) + (uint32_t)((min(lane & 15, NV - 1) * RED_ROW + (lane >> 4) * 16) * 4);
  Fetched nxt = fetch_instance(list, rec, range.x, eff_len - 1 - lane);
  for (int top = eff_len - 1; top >= 0; top -= 32) {
    __syncwarp();  // everyone has left the previous batch
    const int nb = commit_batch(sm, nxt, X0, Y0);
    nxt = fetch_instance(list, rec, range.x, top - 32 - lane);
    for (int j = 0; j < nb; j++) {
      const uint32_t ad = smA + (uint32_t)j * (uint32_t)sizeof(float4);
      const float4 a = lds128(ad);
      const float4 bb = lds128(ad + SM_B);  // cz, thr_m, opacity, position
      const uint32_t fwd = __float_as_uint(bb.w);
      const float dx = __fadd_rn(a.x, -pxf);
      const float qx = __fmul_rn(a.z, dx), hx = __fmul_rn(a.w, dx);
      float dy[BWD_SLOTS], power[BWD_SLOTS];
      bool maybe[BWD_SLOTS];
#pragma unroll
      for (int s = 0; s < BWD_SLOTS; s++) {
        dy[s] = __fadd_rn(a.y, -pyf[s]);
        const float q = __fmaf_rn(dx, qx, __fmul_rn(__fmul_rn(bb.x, dy[s]), dy[s]));
        power[s] = __fmaf_rn(q, -0.5f, -__fmul_rn(hx, dy[s]));
        // entries at or beyond n_contrib were never blended (backward.cu:504-506); power > 0 and the alpha threshold
        // as in the forward.  The pre-test against thr_m avoids expf for pairs that certainly fail alpha >= 1/255.
        maybe[s] = (fwd < lastc[s]) && !(power[s] > 0.0f) && !(power[s] < bb.y);
      }
      bool any_maybe = false;
#pragma unroll
      for (int s = 0; s < BWD_SLOTS; s++) any_maybe |= maybe[s];
      if (!__any_sync(0xffffffffu, any_maybe)) continue;
      float G[BWD_SLOTS], alpha[BWD_SLOTS];
      bool contrib[BWD_SLOTS];
      uint32_t cmask[BWD_SLOTS];
#pragma unroll
      for (int s = 0; s < BWD_SLOTS; s++) {
        G[s] = expf(power[s]);
        alpha[s] = fminf(0.99f, __fmul_rn(bb.z, G[s]));
        contrib[s] = maybe[s] && !(alpha[s] < kAlphaMin);
        cmask[s] = __ballot_sync(0xffffffffu, contrib[s]);
      }
      uint32_t call = 0u;
      bool any_mine = false;
#pragma unroll
      for (int s = 0; s < BWD_SLOTS; s++) { call |= cmask[s]; any_mine |= contrib[s]; }
      if (call == 0u) continue;

      float v[12];
#pragma unroll
      for (int k = 0; k < 12; k++) v[k] = 0.f;
      const float4 c = lds128(ad + SM_C);  // r, g, b, f0
      const float4 d = lds128(ad + SM_D);  // f1, f2, id
#pragma unroll
      for (int s = 0; s < BWD_SLOTS; s++) {
        if (cmask[s] == 0u) continue;  // warp-uniform
        if (contrib[s]) {
          const float one_m_alpha = __fadd_rn(1.f, -alpha[s]);
          // (an IEEE division here, as the reference's T = T / (1 - alpha), was measured: the gradients' distance to
          // the reference does not change — it is set by the reference's own run-to-run atomic-order noise)
#if LSR_BWD_IEEE_DIV
          const float rcp = __fdiv_rn(1.f, one_m_alpha);
          T[s] = __fdiv_rn(T[s], one_m_alpha);  // the reference's T = T / (1 - alpha) (backward.cu:521)
#else
          const float rcp = rcp_ftz(one_m_alpha);  // 1 - alpha is in [0.01, 1]
          T[s] *= rcp;
#endif
          const float w = alpha[s] * T[s];  // d(channel)/d(colour)
          float cd = 0.f;                   // colour . dL_dpixel over all blended channels
          if (DCOL) {
            cd = c.x * dpix[s][0] + c.y * dpix[s][1] + c.z * dpix[s][2];
            v[I_COL + 0] += w * dpix[s][0]; v[I_COL + 1] += w * dpix[s][1]; v[I_COL + 2] += w * dpix[s][2];
          }
          if (FEAT) {
            cd += c.w * dpixF[s][0] + d.x * dpixF[s][1] + d.y * dpixF[s][2];
            v[I_FEAT + 0] += w * dpixF[s][0]; v[I_FEAT + 1] += w * dpixF[s][1]; v[I_FEAT + 2] += w * dpixF[s][2];
          }
          // "colour behind" recurrence on the dot product (backward.cu:527-538, per channel there)
          accA[s] = alpha_last[s] * cd_last[s] + (1.f - alpha_last[s]) * accA[s];
          cd_last[s] = cd;
          alpha_last[s] = alpha[s];
          float dL_dalpha = (cd - accA[s]) * T[s];
          if (DCOL) dL_dalpha += bgT[s] * rcp;  // background term (backward.cu:545-550)
          if (GEOM && LSR_MOMENTS) {
            // moments of the pixel offset about the Gaussian centre, weighted by s = G * dL_dalpha
            const float sw = G[s] * dL_dalpha;
            const float sx = sw * dx, sy = sw * dy[s];
            v[I_OPAC] += sw;
            v[I_MEAN + 0] += sx;
            v[I_MEAN + 1] += sy;
            v[I_CONIC + 0] += sx * dx;
            v[I_CONIC + 1] += sx * dy[s];
            v[I_CONIC + 2] += sy * dy[s];
          } else {
            const float dL_dG = bb.z * dL_dalpha;
            const float gdx = G[s] * dx, gdy = G[s] * dy[s];
            const float dG_ddelx = -gdx * a.z - gdy * a.w;
            const float dG_ddely = -gdy * bb.x - gdx * a.w;
            v[I_MEAN + 0] += dL_dG * dG_ddelx * ddelx_dx;
            v[I_MEAN + 1] += dL_dG * dG_ddely * ddely_dy;
            if (GEOM) {
              v[I_CONIC + 0] += -0.5f * gdx * dx * dL_dG;
              v[I_CONIC + 1] += -0.5f * gdx * dy[s] * dL_dG;
              v[I_CONIC + 2] += -0.5f * gdy * dy[s] * dL_dG;
              v[I_OPAC] += G[s] * dL_dalpha;
            }
          }
        }
      }
      const uint32_t gid = __float_as_uint(d.z);
      if (__popc(call) <= SPARSE_LANES) {
        // very few lanes contribute: their own REDs cost fewer issue slots than the 13-shuffle butterfly
        if (any_mine) {
          atomicAdd(dL_dmean2D + (size_t)gid * 3, v[I_MEAN]);
          atomicAdd(dL_dmean2D + (size_t)gid * 3 + 1, v[I_MEAN + 1]);
          if (GEOM) {
            atomicAdd(dL_dconic + (size_t)gid * 4, v[I_CONIC]);
            atomicAdd(dL_dconic + (size_t)gid * 4 + 1, v[I_CONIC + 1]);
            atomicAdd(dL_dconic + (size_t)gid * 4 + 3, v[I_CONIC + 2]);
            atomicAdd(dL_dopacity + gid, v[I_OPAC]);
          }
          if (DCOL) {
#pragma unroll
            for (int ch = 0; ch < 3; ch++) atomicAdd(dL_dcolors + (size_t)gid * 3 + ch, v[I_COL + ch]);
          }
          if (FEAT) {
#pragma unroll
            for (int ch = 0; ch < 3; ch++) atomicAdd(dL_dfeat + (size_t)gid * 3 + ch, v[I_FEAT + ch]);
          }
        }
      } else {
        if constexpr (SMEM_REDUCE) {
#pragma unroll
          for (int k = 0; k < NV; k++) sts32(red_st + (uint32_t)(k * RED_ROW * 4), v[k]);
          __syncwarp();
          const float4 r0 = lds128(red_ld), r1 = lds128(red_ld + 16), r2 = lds128(red_ld + 32), r3 = lds128(red_ld + 48);
          float t = ((r0.x + r0.y) + (r0.z + r0.w)) + ((r1.x + r1.y) + (r1.z + r1.w)) +
                    (((r2.x + r2.y) + (r2.z + r2.w)) + ((r3.x + r3.y) + (r3.z + r3.w)));
          t += __shfl_xor_sync(0xffffffffu, t, 16);
          if (out_base != nullptr) atomicAdd(out_base + (size_t)gid * out_stride, t);
          __syncwarp();  // the rows are rewritten by the next working instance
        } else {
          multi_reduce<NV, 16>(v, lane);
          if (out_base != nullptr) atomicAdd(out_base + (size_t)gid * out_stride, v[0]);
        }
      }
    }
  }
}

template <bool FEAT, bool DCOL, bool GEOM>
void launch_one(unsigned grid, cudaStream_t s, int W, int H, int tiles_x, const uint2* ranges, const uint32_t* order,
                const uint32_t* list, const GeomRec* rec, const float* bg, const float* final_T, const uint32_t* n_contrib,
                const float* dL_dcolor, const float* dL_dfeature, float* dL_dmean2D, float* dL_dconic,
                float* dL_dopacity, float* dL_dcolors, float* dL_dfeat) {
  render_bwd_kernel<FEAT, DCOL, GEOM><<<grid, 32, 0, s>>>(W, H, tiles_x, ranges, order, list, rec, bg, final_T,
                                                        n_contrib, dL_dcolor, dL_dfeature, dL_dmean2D, dL_dconic, dL_dopacity,
                                                        dL_dcolors, dL_dfeat);
}

}  // namespace

bool blend_acc_supported() {
#if LSR_MOMENTS
  static const bool v1 = getenv("LSR_BWD_V1") != nullptr;
  return !v1;
#else
  return false;
#endif
}

void launch_render_bwd(int W, int H, const uint2* ranges, const uint32_t* order, const uint32_t* list,
                       const GeomRec* rec, const float* bg, bool include_feature, const float* final_T, const uint32_t* n_contrib, const float* dL_dcolor,
                       const float* dL_dfeature, uint32_t need_mask, float* dL_dmean2D, float* dL_dconic,
                       float* dL_dopacity, float* dL_dcolors, float* dL_dfeat, float* acc, cudaStream_t s) {
#if LSR_MOMENTS
  // the two-phase kernel (render_bwd2.cu) is the default; LSR_BWD_V1=1 selects this file's one-phase kernel (A/B)
  if (blend_acc_supported()) {
    launch_render_bwd2(W, H, ranges, order, list, rec, bg, include_feature, final_T, n_contrib, dL_dcolor, dL_dfeature,
                       need_mask, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolors, dL_dfeat, acc, s);
    return;
  }
#endif
  (void)acc;  // the one-phase kernel accumulates into the per-tensor buffers only (lsr_backward passes no record then)
  const int tiles_x = (W + TILE_X - 1) / TILE_X, tiles_y = (H + TILE_Y - 1) / TILE_Y;
  if (tiles_x == 0 || tiles_y == 0) return;
  const unsigned grid = (unsigned)tiles_x * (unsigned)tiles_y * QUADS;
  const bool feat = include_feature && dL_dfeature != nullptr;
  const bool dcol = dL_dcolor != nullptr;
  const bool geom = (need_mask & LSR_NEED_GEOMETRY) != 0;
  if (!feat && !dcol) return;  // no upstream gradient at all: everything stays zero
#define LSR_BWD(F, D, G)                                                                                      \
  launch_one<F, D, G>(grid, s, W, H, tiles_x, ranges, order, list, rec, bg, final_T, n_contrib, dL_dcolor,   \
                      dL_dfeature, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolors, dL_dfeat)
  if (feat) {
    if (dcol) { if (geom) LSR_BWD(true, true, true); else LSR_BWD(true, true, false); }
    else      { if (geom) LSR_BWD(true, false, true); else LSR_BWD(true, false, false); }
  } else {
    if (geom) LSR_BWD(false, true, true); else LSR_BWD(false, true, false);
  }
#undef LSR_BWD
  count_launch();
}

}  // namespace lsr
