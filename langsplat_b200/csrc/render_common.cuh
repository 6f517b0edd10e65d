// render_common.cuh — batch staging shared by the forward and backward blend kernels.
//
// Per 256-instance batch every thread gathers one GeomRec and decides, once for the whole tile, whether that Gaussian
// can reach alpha >= 1/255 at ANY pixel of the tile (exact minimum of the conic quadratic form over the tile rectangle
// against the per-Gaussian threshold ln(1/(255*opacity))).  Instances that cannot are dropped before the per-pixel
// loop; the survivors are compacted in order, each keeping its original position in the tile's list so n_contrib and
// the backward's `contributor` bookkeeping are unchanged.  The reference examines every instance at every pixel
// (forward.cu:330-372, backward.cu:498-522); dropping provably non-contributing ones changes no result bit.
//
// The same threshold gives a cheap per-pixel pre-test: power < thr - eps  =>  opacity*exp(power) < 1/255, so expf is
// only evaluated for pairs that can pass.  eps (>= 1e-5) covers expf's <= 2 ulp error and the product rounding, so a
// pair the reference accepts is never pre-rejected; pairs inside the eps band take the exact path.
#pragma once
#include "common.cuh"

namespace lsr {

template <int NT>  // NT = threads per CTA = instances staged per batch
struct BatchSmemT {
  float4 A[NT];  // x, y, conic.x, conic.y
  float4 B[NT];  // conic.z, thr_m (pre-test threshold), opacity, position in the tile list (uint bits)
  float4 C[NT];  // r, g, b, f0
  float4 D[NT];  // f1, f2, gaussian id (uint bits), unused
  uint32_t wcnt[NT / 32];
};
using BatchSmem = BatchSmemT<TILE_PIX>;

constexpr float kAlphaMin = 1.0f / 255.0f;

// true if the Gaussian may reach alpha >= 1/255 somewhere in the 16x16 pixel block whose first pixel is (X0, Y0)
__device__ __forceinline__ bool tile_may_contribute(float mx, float my, float a, float b, float c, float thr,
                                                    float X0, float Y0) {
  const float dx_hi = mx - X0, dx_lo = dx_hi - (float)(TILE_X - 1);
  const float dy_hi = my - Y0, dy_lo = dy_hi - (float)(TILE_Y - 1);
  if (!(a > 0.f && c > 0.f && a * c - b * b > 0.f)) return true;  // not positive definite (or NaN): no claim
  if (dx_lo <= 0.f && dx_hi >= 0.f && dy_lo <= 0.f && dy_hi >= 0.f) return true;  // centre inside the block
  const float inv_c = __fdividef(1.f, c), inv_a = __fdividef(1.f, a);
  float qmin;
  {
    float t = fminf(fmaxf(-b * dx_lo * inv_c, dy_lo), dy_hi);
    qmin = a * dx_lo * dx_lo + 2.f * b * dx_lo * t + c * t * t;
    t = fminf(fmaxf(-b * dx_hi * inv_c, dy_lo), dy_hi);
    qmin = fminf(qmin, a * dx_hi * dx_hi + 2.f * b * dx_hi * t + c * t * t);
    t = fminf(fmaxf(-b * dy_lo * inv_a, dx_lo), dx_hi);
    qmin = fminf(qmin, a * t * t + 2.f * b * t * dy_lo + c * dy_lo * dy_lo);
    t = fminf(fmaxf(-b * dy_hi * inv_a, dx_lo), dx_hi);
    qmin = fminf(qmin, a * t * t + 2.f * b * t * dy_hi + c * dy_hi * dy_hi);
  }
  const float dxm = fmaxf(fabsf(dx_lo), fabsf(dx_hi)), dym = fmaxf(fabsf(dy_lo), fabsf(dy_hi));
  const float smax = a * dxm * dxm + c * dym * dym + 2.f * fabsf(b) * dxm * dym;
  // margin: rounding of the reference's own power evaluation (a few ulp of its largest term) plus ours
  const float margin = 1e-5f + 2e-6f * smax;
  return !(-0.5f * qmin + margin < thr);
}

// Stage one batch.  `list_index` is the index into the tile's range (0-based from range.x) this thread fetches, or -1.
// Returns the number of surviving entries (identical in all threads).  Must be called by all NT threads of the CTA.
template <int NT>
__device__ __forceinline__ int stage_batch(BatchSmemT<NT>& sm, const uint32_t* __restrict__ list,
                                           const GeomRec* __restrict__ rec, uint32_t range_x, int list_index,
                                           float X0, float Y0) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a, c = a;
  uint32_t id = 0;
  bool keep = false;
  float thr_m = 0.f;
  if (list_index >= 0) {
    id = __ldg(list + range_x + list_index);
    const float4* r4 = reinterpret_cast<const float4*>(rec + id);
    a = __ldg(r4);
    b = __ldg(r4 + 1);
    c = __ldg(r4 + 2);
    const float o = b.y;
    // alpha = min(0.99, o*exp(p)) < 1/255  <=>  p < ln(1/(255 o))   (o > 0).  o <= 0 never passes; NaN always does.
    const float thr = (o > 0.f) ? __logf(__fdividef(kAlphaMin, o)) : ((o == o) ? __int_as_float(0x7f800000)
                                                                              : __int_as_float(0xff800000));
    thr_m = thr - (2e-5f + 2e-6f * fabsf(thr));
    keep = tile_may_contribute(a.x, a.y, a.z, a.w, b.x, thr_m, X0, Y0);
  }
  const uint32_t m = __ballot_sync(0xffffffffu, keep);
  if (lane == 0) sm.wcnt[warp] = __popc(m);
  __syncthreads();
  int base = 0, total = 0;
#pragma unroll
  for (int w = 0; w < NT / 32; w++) {
    const int n = (int)sm.wcnt[w];
    if (w < warp) base += n;
    total += n;
  }
  if (keep) {
    const int slot = base + __popc(m & ((1u << lane) - 1));
    sm.A[slot] = a;                                                              // x, y, cx, cy
    sm.B[slot] = make_float4(b.x, thr_m, b.y, __uint_as_float((uint32_t)list_index));  // cz, thr_m, o, pos
    sm.C[slot] = make_float4(b.z, b.w, c.x, c.y);                                // r, g, b, f0
    sm.D[slot] = make_float4(c.z, c.w, __uint_as_float(id), 0.f);                // f1, f2, id
  }
  __syncthreads();
  return total;
}

}  // namespace lsr
