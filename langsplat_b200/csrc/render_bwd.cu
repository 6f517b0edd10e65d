// render_bwd.cu — K7': per-pixel reverse blend; gradients w.r.t. mean2D, conic, opacity, colour, language feature.
// Replaces BACKWARD::render / renderCUDA (RAST/cuda_rasterizer/backward.cu:399-593).
//
// The reference issues 12 global atomicAdds per contributing (pixel, Gaussian) pair (backward.cu:541-590).  Here all
// 32 lanes of a warp walk the same instance in lock-step, so the (up to) 12 per-lane contributions are first summed
// across the warp with a value-halving butterfly (13 shuffles for 12 values instead of 60: each round a lane keeps half
// of its values and sends the other half to its partner), leaving each value's warp total on one lane, which issues a
// single RED.ADD.F32.  Instances no lane contributes to are dropped after one ballot.
// The walk starts at the CTA-wide maximum of n_contrib, so the never-blended tail of a tile's list is not even fetched.
// Template switches specialise on what the caller needs (ctx.needs_input_grad): FEAT = language feature blended,
// DCOL = a colour upstream gradient exists, GEOM = conic/opacity gradients wanted.  LangSplat's feature-only training
// (train.py:96-99 with gaussian_model.py:212-217) runs <FEAT=1, DCOL=0, GEOM=0>: 5 values instead of 12.
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

template <bool FEAT, bool DCOL, bool GEOM>
__global__ void __launch_bounds__(TILE_PIX)
render_bwd_kernel(int W, int H, const uint2* __restrict__ ranges, const uint32_t* __restrict__ list,
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

  __shared__ BatchSmem sm;
  __shared__ uint32_t s_max[TILE_PIX / 32];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // warp w owns an 8x4-pixel footprint (compact, instead of the reference's 16x2 strips)
  const int tx = (warp & 1) * 8 + (lane & 7), ty = (warp >> 1) * 4 + (lane >> 3);
  const int tiles_x = (W + TILE_X - 1) / TILE_X;
  const uint32_t px = blockIdx.x * TILE_X + tx, py = blockIdx.y * TILE_Y + ty;
  const bool inside = px < (uint32_t)W && py < (uint32_t)H;
  const float pxf = (float)px, pyf = (float)py;
  const float X0 = (float)(blockIdx.x * TILE_X), Y0 = (float)(blockIdx.y * TILE_Y);
  const size_t N = (size_t)W * H;
  const size_t pix = (size_t)py * W + px;
  const uint2 range = ranges[blockIdx.y * tiles_x + blockIdx.x];

  const float T_final = inside ? final_Ts[pix] : 0.f;
  float T = T_final;
  const uint32_t last_contributor = inside ? n_contrib[pix] : 0u;

  // The CTA-wide maximum of n_contrib bounds the part of the list anyone needs.
  uint32_t wmax = last_contributor;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, d));
  if (lane == 0) s_max[warp] = wmax;
  __syncthreads();
  uint32_t bmax = 0;
#pragma unroll
  for (int i = 0; i < TILE_PIX / 32; i++) bmax = max(bmax, s_max[i]);
  const int eff_len = (int)bmax;  // forward indices [0, eff_len) are the only ones examined

  float dpix[3] = {0.f, 0.f, 0.f}, dpixF[3] = {0.f, 0.f, 0.f};
  if (inside) {
    if (DCOL) {
#pragma unroll
      for (int i = 0; i < 3; i++) dpix[i] = dL_dpixels[i * N + pix];
    }
    if (FEAT) {
#pragma unroll
      for (int i = 0; i < 3; i++) dpixF[i] = dL_dpixels_F[i * N + pix];
    }
  }
  float bg_dot = 0.f;
  if (DCOL) bg_dot = __ldg(bg) * dpix[0] + __ldg(bg + 1) * dpix[1] + __ldg(bg + 2) * dpix[2];

  float accum_rec[3] = {0.f, 0.f, 0.f}, last_color[3] = {0.f, 0.f, 0.f};
  float accum_recF[3] = {0.f, 0.f, 0.f}, last_F[3] = {0.f, 0.f, 0.f};
  float last_alpha = 0.f;
  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;

  // which reduced value this lane ends up holding, and where it goes
  const int held = held_index<NV, 16>(lane);
  float* out_base = nullptr;
  int out_stride = 0;
  if (held >= 0) {
    if (held < I_MEAN + 2) { out_base = dL_dmean2D + held; out_stride = 3; }
    else if (GEOM && held < I_OPAC) { const int k = held - I_CONIC; out_base = dL_dconic + (k == 2 ? 3 : k); out_stride = 4; }
    else if (GEOM && held == I_OPAC) { out_base = dL_dopacity; out_stride = 1; }
    else if (DCOL && held >= I_COL && held < I_COL + 3) { out_base = dL_dcolors + (held - I_COL); out_stride = 3; }
    else if (FEAT && held >= I_FEAT) { out_base = dL_dfeat + (held - I_FEAT); out_stride = 3; }
  }

  const int rounds = (eff_len + TILE_PIX - 1) / TILE_PIX;
  for (int b = 0; b < rounds; b++) {
    // back to front: this batch covers forward indices eff_len-1-b*256 downwards (backward.cu:486-498)
    const int fwd_fetch = eff_len - 1 - (b * TILE_PIX + tid);
    const int nb = stage_batch(sm, list, rec, range.x, fwd_fetch >= 0 ? fwd_fetch : -1, X0, Y0);
    for (int j = 0; j < nb; j++) {
      const float4 a = sm.A[j];
      const float4 bb = sm.B[j];  // cz, thr_m, opacity, position
      const uint32_t fwd = __float_as_uint(bb.w);
      const float dx = __fadd_rn(a.x, -pxf), dy = __fadd_rn(a.y, -pyf);
      const float q = __fmaf_rn(dx, __fmul_rn(a.z, dx), __fmul_rn(__fmul_rn(bb.x, dy), dy));
      const float power = __fmaf_rn(q, -0.5f, -__fmul_rn(__fmul_rn(a.w, dx), dy));
      // entries at or beyond n_contrib were never blended (backward.cu:504-506); power > 0 and the alpha threshold
      // as in the forward.  The pre-test against thr_m avoids expf for pairs that certainly fail alpha >= 1/255.
      const bool maybe = (fwd < last_contributor) && !(power > 0.0f) && !(power < bb.y);
      if (__ballot_sync(0xffffffffu, maybe) == 0u) continue;
      const float G = expf(power);
      const float alpha = fminf(0.99f, __fmul_rn(bb.z, G));
      const bool contrib = maybe && !(alpha < kAlphaMin);
      const uint32_t cmask = __ballot_sync(0xffffffffu, contrib);
      if (cmask == 0u) continue;

      float v[12];
#pragma unroll
      for (int k = 0; k < 12; k++) v[k] = 0.f;
      const float4 c = sm.C[j];  // r, g, b, f0
      const float4 d = sm.D[j];  // f1, f2, id
      if (contrib) {
        const float one_m_alpha = __fadd_rn(1.f, -alpha);
        T = __fdiv_rn(T, one_m_alpha);
        const float dchannel_dcolor = alpha * T;
        float dL_dalpha = 0.0f;
        if (DCOL) {
          const float col[3] = {c.x, c.y, c.z};
#pragma unroll
          for (int ch = 0; ch < 3; ch++) {
            accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
            last_color[ch] = col[ch];
            dL_dalpha += (col[ch] - accum_rec[ch]) * dpix[ch];
            v[I_COL + ch] = dchannel_dcolor * dpix[ch];
          }
        }
        if (FEAT) {
          const float ft[3] = {c.w, d.x, d.y};
#pragma unroll
          for (int ch = 0; ch < 3; ch++) {
            accum_recF[ch] = last_alpha * last_F[ch] + (1.f - last_alpha) * accum_recF[ch];
            last_F[ch] = ft[ch];
            dL_dalpha += (ft[ch] - accum_recF[ch]) * dpixF[ch];
            v[I_FEAT + ch] = dchannel_dcolor * dpixF[ch];
          }
        }
        dL_dalpha *= T;
        last_alpha = alpha;
        if (DCOL) {
          if (bg_dot != 0.f) dL_dalpha += (-T_final / one_m_alpha) * bg_dot;
        }
        const float dL_dG = bb.z * dL_dalpha;
        const float gdx = G * dx, gdy = G * dy;
        const float dG_ddelx = -gdx * a.z - gdy * a.w;
        const float dG_ddely = -gdy * bb.x - gdx * a.w;
        v[I_MEAN + 0] = dL_dG * dG_ddelx * ddelx_dx;
        v[I_MEAN + 1] = dL_dG * dG_ddely * ddely_dy;
        if (GEOM) {
          v[I_CONIC + 0] = -0.5f * gdx * dx * dL_dG;
          v[I_CONIC + 1] = -0.5f * gdx * dy * dL_dG;
          v[I_CONIC + 2] = -0.5f * gdy * dy * dL_dG;
          v[I_OPAC] = G * dL_dalpha;
        }
      }
      const uint32_t gid = __float_as_uint(d.z);
      if (__popc(cmask) <= SPARSE_LANES) {
        // very few lanes contribute: their own REDs cost fewer issue slots than the 13-shuffle butterfly
        if (contrib) {
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
        multi_reduce<NV, 16>(v, lane);
        if (out_base != nullptr) atomicAdd(out_base + (size_t)gid * out_stride, v[0]);
      }
    }
  }
}

template <bool FEAT, bool DCOL, bool GEOM>
void launch_one(dim3 grid, cudaStream_t s, int W, int H, const uint2* ranges, const uint32_t* list, const GeomRec* rec,
                const float* bg, const float* final_T, const uint32_t* n_contrib, const float* dL_dcolor,
                const float* dL_dfeature, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolors,
                float* dL_dfeat) {
  render_bwd_kernel<FEAT, DCOL, GEOM><<<grid, TILE_PIX, 0, s>>>(W, H, ranges, list, rec, bg, final_T, n_contrib,
                                                                 dL_dcolor, dL_dfeature, dL_dmean2D, dL_dconic,
                                                                 dL_dopacity, dL_dcolors, dL_dfeat);
}

}  // namespace

void launch_render_bwd(int W, int H, const uint2* ranges, const uint32_t* list, const GeomRec* rec, const float* bg,
                       bool include_feature, const float* final_T, const uint32_t* n_contrib, const float* dL_dcolor,
                       const float* dL_dfeature, uint32_t need_mask, float* dL_dmean2D, float* dL_dconic,
                       float* dL_dopacity, float* dL_dcolors, float* dL_dfeat, cudaStream_t s) {
  dim3 grid((W + TILE_X - 1) / TILE_X, (H + TILE_Y - 1) / TILE_Y);
  if (grid.x == 0 || grid.y == 0) return;
  const bool feat = include_feature && dL_dfeature != nullptr;
  const bool dcol = dL_dcolor != nullptr;
  const bool geom = (need_mask & LSR_NEED_GEOMETRY) != 0;
  if (!feat && !dcol) return;  // no upstream gradient at all: everything stays zero
#define LSR_BWD(F, D, G)                                                                                           \
  launch_one<F, D, G>(grid, s, W, H, ranges, list, rec, bg, final_T, n_contrib, dL_dcolor, dL_dfeature, dL_dmean2D, \
                      dL_dconic, dL_dopacity, dL_dcolors, dL_dfeat)
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
