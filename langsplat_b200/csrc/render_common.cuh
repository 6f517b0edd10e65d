// render_common.cuh — instance staging shared by the forward and backward blend kernels.
//
// The unit of work is one WARP = one 8x8-pixel quadrant of a 16x16 tile, two pixels per thread (rows r and r+4);
// a CTA is a single warp, so nothing on the blend path needs a CTA barrier and a quadrant retires as soon as its own
// 64 pixels are done (the reference couples all 256 pixels of a tile through __syncthreads, forward.cu:313-372).
// The four quadrants of a tile read the same tile range; the records they share come out of L2.
//
// Per 32-instance batch every lane gathers one GeomRec and decides whether that Gaussian can reach alpha >= 1/255 at
// ANY pixel of the quadrant (exact minimum of the conic quadratic form over the quadrant's pixel rectangle against the
// per-Gaussian threshold ln(1/(255*opacity))).  Instances that cannot are dropped before the per-pixel loop; the
// survivors are compacted in order (ballot + popc), each keeping its original position in the tile's list so n_contrib
// and the backward's `contributor` bookkeeping are unchanged.  The reference examines every instance at every pixel
// (forward.cu:330-372, backward.cu:498-522); dropping provably non-contributing ones changes no result bit.
// (Bench scene: 6.2 M (quadrant, instance) pairs with a tile-level test only, 3.7 M with the per-quadrant test, of
// which 3.6 M have at least one contributing pixel.)
//
// The same threshold gives a cheap per-pixel pre-test: power < thr - eps  =>  opacity*exp(power) < 1/255, so expf is
// only evaluated for pairs that can pass.  eps (>= 1e-5) covers expf's <= 2 ulp error and the product rounding, so a
// pair the reference accepts is never pre-rejected; pairs inside the eps band take the exact path.
//
// Staging is split into fetch (global loads into registers) and commit (test + shared-memory writes) so that the
// kernels issue the next batch's loads before they blend the current one.  (A cp.async variant that parks the next
// batch in shared memory instead of 13 registers was measured: 64 registers and 32 warps per SM, but render_bwd went
// from 852 to 875 us — the extra wait + LDS round trip per batch costs more than the occupancy buys.)
#pragma once
#include <type_traits>

#include "common.cuh"

namespace lsr {

constexpr int QUAD_X = 8, QUAD_Y = 8;  // pixel footprint of one warp
constexpr int QUADS_X = TILE_X / QUAD_X, QUADS_Y = TILE_Y / QUAD_Y;
constexpr int QUADS = QUADS_X * QUADS_Y;  // warps (= CTAs) per tile
constexpr int BLEND_SLOTS = 2;            // pixels per thread (rows r and r+4)
static_assert(QUAD_X * QUAD_Y == 32 * BLEND_SLOTS, "a warp covers its quadrant with BLEND_SLOTS pixels per lane");

struct WarpBatch {
  float4 A[32];  // x, y, conic.x, conic.y
  float4 B[32];  // conic.z, thr_m (pre-test threshold), opacity, position in the tile list (uint bits)
  float4 C[32];  // r, g, b, f0
  float4 D[32];  // f1, f2, gaussian id (uint bits), unused
};

// Packed layout for the forward blend: sm_100 issues two independent IEEE fp32 FMAs as one instruction
// (fma.rn.f32x2 -> FFMA2).  A thread's two pixels share x and every per-Gaussian constant, so the constants are stored
// already duplicated: each 64-bit half of a float4 is directly a packed operand and no broadcast MOVs are needed.
struct WarpBatchPk {
  float4 A[32];  // x, x, y, y
  float4 Q[32];  // conic.x, conic.x, -conic.y, -conic.y
  float4 B[32];  // conic.z, conic.z, thr_m, opacity
  float4 C[32];  // r, g, b, f0
  float4 D[32];  // f1, f2, position in the tile list (uint bits), gaussian id (uint bits)
};

typedef unsigned long long f32x2;  // two floats in one 64-bit register pair: lo = slot 0, hi = slot 1
__device__ __forceinline__ f32x2 pk2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void upk2(f32x2 v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}

constexpr float kAlphaMin = 1.0f / 255.0f;

// MUFU approximations without the denormal fix-up code (inputs flushed to zero, which every use below tolerates)
__device__ __forceinline__ float rcp_ftz(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lg2_ftz(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// explicit shared-window loads: the address is formed once per kernel instead of once per access
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ float4 lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(addr)
               : "memory");
  return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
constexpr uint32_t SM_B = 32 * sizeof(float4), SM_C = 2 * SM_B, SM_D = 3 * SM_B;  // offsets of B, C, D from A

// true if the Gaussian may reach alpha >= 1/255 somewhere in the pixel rectangle [X0, X0+ex] x [Y0, Y0+ey]
__device__ __forceinline__ bool rect_may_contribute(float mx, float my, float a, float b, float c, float thr,
                                                    float X0, float Y0, float ex, float ey) {
  const float dx_hi = mx - X0, dx_lo = dx_hi - ex;
  const float dy_hi = my - Y0, dy_lo = dy_hi - ey;
  if (!(a > 0.f && c > 0.f && a * c - b * b > 0.f)) return true;  // not positive definite (or NaN): no claim
  if (dx_lo <= 0.f && dx_hi >= 0.f && dy_lo <= 0.f && dy_hi >= 0.f) return true;  // centre inside the rectangle
  const float inv_c = rcp_ftz(c), inv_a = rcp_ftz(a);
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

// One instance as fetched from global memory (registers).
struct Fetched {
  float4 a, b, c;  // the GeomRec
  uint32_t id;
  int pos;         // position in the tile's list, -1 = nothing fetched
};

__device__ __forceinline__ Fetched fetch_instance(const uint32_t* __restrict__ list, const GeomRec* __restrict__ rec,
                                                  uint32_t range_x, int list_index) {
  Fetched f;
  f.a = make_float4(0.f, 0.f, 0.f, 0.f);
  f.b = f.a;
  f.c = f.a;
  f.id = 0;
  f.pos = list_index;
  if (list_index >= 0) {
    f.id = __ldg(list + range_x + list_index);
    const float4* r4 = reinterpret_cast<const float4*>(rec + f.id);
    f.a = __ldg(r4);
    f.b = __ldg(r4 + 1);
    f.c = __ldg(r4 + 2);
  }
  return f;
}

// Tests the warp's 32 fetched instances against the quadrant whose first pixel is (X0, Y0) and writes the survivors
// to shared memory in lane order.  Returns their number (identical in all lanes).  All 32 lanes must call it, after a
// __syncwarp() that ends the reads of the previous batch; it ends with a __syncwarp().
template <typename Batch>
__device__ __forceinline__ int commit_batch(Batch& sm, const Fetched& f, float X0, float Y0) {
  const int lane = threadIdx.x & 31;
  bool keep = false;
  float thr_m = 0.f;
  if (f.pos >= 0) {
    const float o = f.b.y;
    // alpha = min(0.99, o*exp(p)) < 1/255  <=>  p < ln(1/(255 o))   (o > 0).  o <= 0 never passes; NaN always does.
    // ln(1/(255 o)) = (log2(1/255) - log2(o)) * ln 2; a denormal o is flushed to 0 -> thr = +inf, and indeed
    // o*exp(p) <= o < 1/255 then.
    const float thr = (o > 0.f) ? (-7.99435344f - lg2_ftz(o)) * 0.693147181f
                                : ((o == o) ? __int_as_float(0x7f800000) : __int_as_float(0xff800000));
    thr_m = thr - (2e-5f + 2e-6f * fabsf(thr));
    keep = rect_may_contribute(f.a.x, f.a.y, f.a.z, f.a.w, f.b.x, thr_m, X0, Y0, (float)(QUAD_X - 1),
                               (float)(QUAD_Y - 1));
  }
  const uint32_t m = __ballot_sync(0xffffffffu, keep);
  if (keep) {
    const int slot = __popc(m & ((1u << lane) - 1u));
    if constexpr (std::is_same<Batch, WarpBatchPk>::value) {
      sm.A[slot] = make_float4(f.a.x, f.a.x, f.a.y, f.a.y);
      sm.Q[slot] = make_float4(f.a.z, f.a.z, -f.a.w, -f.a.w);
      sm.B[slot] = make_float4(f.b.x, f.b.x, thr_m, f.b.y);
      sm.C[slot] = make_float4(f.b.z, f.b.w, f.c.x, f.c.y);
      sm.D[slot] = make_float4(f.c.z, f.c.w, __uint_as_float((uint32_t)f.pos), __uint_as_float(f.id));
    } else {
      sm.A[slot] = f.a;                                                                // x, y, cx, cy
      sm.B[slot] = make_float4(f.b.x, thr_m, f.b.y, __uint_as_float((uint32_t)f.pos));  // cz, thr_m, o, pos
      sm.C[slot] = make_float4(f.b.z, f.b.w, f.c.x, f.c.y);                            // r, g, b, f0
      sm.D[slot] = make_float4(f.c.z, f.c.w, __uint_as_float(f.id), 0.f);              // f1, f2, id
    }
  }
  __syncwarp();
  return __popc(m);
}

}  // namespace lsr
