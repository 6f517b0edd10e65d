// render_bwd2.cu — K7'' two-phase reverse blend (selected by default; LSR_BWD_V1=1 keeps render_bwd.cu's kernel for A/B).
// Replaces BACKWARD::render / renderCUDA (RAST/cuda_rasterizer/backward.cu:399-593), same contract as render_bwd.cu.
//
// Why.  ncu on the one-phase kernel (profiles/r2a): 203 warp instructions per working (quadrant, instance) pair, of
// which 55 are the cross-lane reduction of the 12 per-lane partial sums (12 STS + 4 LDS.128 + 15 FADD + SHFL + RED) and
// 30 the multiply-adds that form those partials; 38 shared-memory wavefronts per pair, l1tex 80 % busy next to an
// 82 % busy issue stage.  Every gradient of a pair is a product of ONE of two per-(pixel, instance) scalars with
// something that depends on the pixel or on the instance alone:
//     w = alpha * T                (weights dL_dcolour, dL_dfeature:  sum over pixels of w * dL_dpixel[ch])
//     s = G * dL_dalpha            (weights the six moments M0, Mx, My, Mxx, Mxy, Myy of the pixel offset, which the
//                                   per-Gaussian kernel turns into dL_dopacity / dL_dmean2D / dL_dconic)
// so the walk only has to produce (w, s), and the twelve sums can be formed with the instance, not the pixel, as the
// parallel index — where they are plain per-lane accumulations and nothing has to be reduced across the warp.
//
// Phase A (lane = 2 pixels of the 8x8 quadrant, as in the forward): the reverse walk with its two sequential
//   recurrences (T, colour-behind), identical decisions; per working instance it stores w and s of its two pixels as
//   one COLUMN of a [8 instances][64 pixels] shared-memory panel (2 STS per pixel slot, conflict-free).
// Phase B (every 8 working instances; lane = (instance c, quarter q of the pixels)): each lane runs over its 16 pixels
//   of its column, four at a time (LDS.128 of w, of s, and of the per-pixel constants dL_dpixel[6], -x, -y, which sit
//   channel-major in shared memory, written once per quadrant), accumulating with packed fp32 instructions
//   (fma.rn.f32x2 over two pixels at once).  Two value-halving shuffle rounds join the four quarters and leave each
//   lane with three finished sums of one tensor (q = 0: Mx, My, M0; 1: Mxx, Mxy, Myy; 2: colour; 3: feature), which
//   it adds to global memory with three REDs: 12 REDs per instance as before, but issued as 3 full-width instructions
//   per 8 instances instead of 8 mostly-empty ones.
// Per working pair: ~117 (phase A) + ~30 (phase B) instructions instead of 203, and ~20 shared wavefronts instead of 38.
// Sums are formed in a different order than in render_bwd.cu (and than the reference's atomics): gradients carry the
// path's 1e-4 relative tolerance; the forward decisions are untouched.
#include "render_common.cuh"

#if LSR_MOMENTS  // the two-phase backward hands moments to the per-Gaussian kernel (an LSR_MOMENTS=0 build uses render_bwd.cu)

namespace lsr {

namespace {

__device__ __forceinline__ float bwd_exp(float x) {
#if LSR_BWD2_FAST_EXP
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x * 1.4426950408889634f));
  return y;
#else
  return expf(x);
#endif
}
constexpr float kAlphaBand = 4e-6f / 255.0f;  // LSR_BWD2_FAST_EXP == 2: |alpha - 1/255| below this -> expf decides

constexpr int NCOL = 8;                // working instances per phase-B pass
constexpr int PARTS = 32 / NCOL;       // lanes per instance in phase B
constexpr int PART_PIX = 64 / PARTS;   // pixels per lane in phase B (16)
constexpr int ROW = 68;                // floats per panel row: pixel p sits at p + 4 * (p >> 5) (bank spreading)
constexpr int PC_ROWS = 8;             // per-pixel constants, channel-major: dpix[3], dpixF[3], -x, -y
static_assert(NCOL == 8 && PARTS == 4 && PART_PIX == 16, "the lane mapping below is written for 8 x 4");

struct __align__(16) Bwd2Smem {
  WarpBatch batch;            // staged instance records (render_common.cuh)
  float W[NCOL][ROW];         // w = alpha * T per (instance, pixel)
  float S[NCOL][ROW];         // s = G * dL_dalpha
  float PC[PC_ROWS][ROW];     // per-pixel constants
  float4 meta[NCOL];          // centre x, y, Gaussian id (uint bits), opacity
  float4 conic[NCOL];         // conic x, y, z (only read when the mean gradient is finished here: !GEOM)
};

#ifndef LSR_BWD2_MINB
#define LSR_BWD2_MINB 24
#endif
// Chain shorteners for the default walk (A/B switches): request the colour record with the geometry record, one
// warp-wide vote per instance instead of three, read the next instance's record one iteration ahead.
#ifndef LSR_BWD2_EARLY_CD
#define LSR_BWD2_EARLY_CD 0
#endif
#ifndef LSR_BWD2_ONE_VOTE
// Measured (GPU call r2m, C2 / C3): three votes 826 / 538 us, one vote 802 / 520 us — every warp-wide vote is a stop
// in the walk's dependent chain.  (Early colour record: no change; record prefetch: 937 us, it spills.)
// 2: no vote at all — an instance nobody contributes to (rare after the batch's rectangle test) takes a column of
//    zeros.  GPU call r2n: 804 -> 802 us at C2, 522 -> 493 us at C3.
#define LSR_BWD2_ONE_VOTE 2
#endif
#ifndef LSR_BWD2_PREFETCH
#define LSR_BWD2_PREFETCH 0
#endif
#ifndef LSR_BWD2_FAST_EXP
// 1: G = ex2.approx(power * log2 e) (2 instructions) instead of expf (8: range reduction around the same MUFU.EX2).
//    Both are accurate to ~2 ulp, but they are different functions: a pair whose alpha sits within an ulp of 1/255 may
//    be kept here and dropped by the forward (or the reverse), which moves one pixel's T by 0.4 %.  Measured (GPU call
//    r2n): 804 -> 774 us at C2, but the full-size C3 / C4 parity tests fail (means2D 2.7e-4 of the largest element).
// 2: as 1, but a pair whose alpha lands within 4e-6 (relative) of 1/255 is re-evaluated with expf, so every keep /
//    drop decision is the forward's (both functions are within 1e-6 of exp); G itself differs by <= 4e-7 relative.
//    Parity green, but SLOWER than expf (GPU call r2o: 819 against 800 us at C2, 529 against 492 us at C3): the band
//    test and its branch cost more than the six instructions saved.  expf stays.
#define LSR_BWD2_FAST_EXP 0
#endif
#ifndef LSR_BWD2_PIPE
// 0: one instance at a time, early-outs, per-slot branches (the default).
// 1: software-pipelined, branch-free walk: front half of instance j+1 next to the back half of instance j.  Measured
//    (GPU call r2d, C2): 945 us against 826 us — the extra work (exp for every examined pair, selects, the re-read
//    records) and the registers cost more than the overlap returns; with a 96-register cap 868 us.
// 2: as 0, but the back half of an instance is branch-free, so that the two pixel slots' chains share a basic block.
#define LSR_BWD2_PIPE 0
#endif

template <bool FEAT, bool DCOL, bool GEOM>
__global__ void __launch_bounds__(32, LSR_BWD2_MINB)
render_bwd2_kernel(int W, int H, int tiles_x, const uint2* __restrict__ ranges, const uint32_t* __restrict__ order,
                   const uint32_t* __restrict__ list, const GeomRec* __restrict__ rec, const float* __restrict__ bg,
                   const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                   const float* __restrict__ dL_dpixels, const float* __restrict__ dL_dpixels_F,
                   float* __restrict__ dL_dmean2D, float* __restrict__ dL_dconic, float* __restrict__ dL_dopacity,
                   float* __restrict__ dL_dcolors, float* __restrict__ dL_dfeat, float* __restrict__ acc) {
  __shared__ Bwd2Smem sm;

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

  float pyf[BLEND_SLOTS], T[BLEND_SLOTS], bgT[BLEND_SLOTS];
  uint32_t lastc[BLEND_SLOTS];
  float dpix[BLEND_SLOTS][3], dpixF[BLEND_SLOTS][3];
  float accA[BLEND_SLOTS], cd_last[BLEND_SLOTS], alpha_last[BLEND_SLOTS];
  uint32_t wmax = 0;
  const uint32_t smPC = smem_u32(&sm.PC[0][0]);
#pragma unroll
  for (int s = 0; s < BLEND_SLOTS; s++) {
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
    // this pixel's constants for phase B: pixel p = 32 s + lane of the quadrant sits at column p + 4 (p >> 5)
    const uint32_t col = (uint32_t)(lane + 36 * s) * 4u;
#pragma unroll
    for (int i = 0; i < 3; i++) {
      if (DCOL) sts32(smPC + (uint32_t)(i * ROW * 4) + col, dpix[s][i]);
      if (FEAT) sts32(smPC + (uint32_t)((3 + i) * ROW * 4) + col, dpixF[s][i]);
    }
    sts32(smPC + (uint32_t)(6 * ROW * 4) + col, -pxf);
    sts32(smPC + (uint32_t)(7 * ROW * 4) + col, -pyf[s]);
  }
  // The quadrant's maximum of n_contrib bounds the part of the list anyone here needs.
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, d));
  const int eff_len = (int)wmax;  // forward indices [0, eff_len) are the only ones examined

  // ---- phase B: the panel's columns -> twelve sums per instance -> global memory ----
  const int pb_c = lane & (NCOL - 1), pb_q = lane >> 3;
  const uint32_t pb_off = (uint32_t)(PART_PIX * pb_q + 4 * (pb_q >> 1)) * 4u;  // first pixel of this lane's quarter
  const uint32_t pb_w = smem_u32(&sm.W[0][0]) + (uint32_t)(pb_c * ROW * 4) + pb_off;
  const uint32_t pb_s = smem_u32(&sm.S[0][0]) + (uint32_t)(pb_c * ROW * 4) + pb_off;
  const uint32_t pb_pc = smPC + pb_off;
  const uint32_t pb_meta = smem_u32(&sm.meta[0]) + (uint32_t)pb_c * 16u;
  const uint32_t pb_conic = smem_u32(&sm.conic[0]) + (uint32_t)pb_c * 16u;
  const float half_W = 0.5f * W, half_H = 0.5f * H;

  auto flush = [&](int n_used) {
    const float4 m = lds128(pb_meta);  // centre x, y, id, opacity
    const f32x2 gx2 = pk2(m.x, m.x), gy2 = pk2(m.y, m.y);
    const f32x2 zero2 = pk2(0.f, 0.f);
    f32x2 aM0 = zero2, aMx = zero2, aMy = zero2, aMxx = zero2, aMxy = zero2, aMyy = zero2;
    f32x2 aC[3] = {zero2, zero2, zero2}, aF[3] = {zero2, zero2, zero2};
#pragma unroll
    for (int k = 0; k < PART_PIX / 4; k++) {
      const uint32_t o = (uint32_t)k * 16u;
      const float4 w4 = lds128(pb_w + o), s4 = lds128(pb_s + o);
      const float4 nx = lds128(pb_pc + (uint32_t)(6 * ROW * 4) + o), ny = lds128(pb_pc + (uint32_t)(7 * ROW * 4) + o);
      const f32x2 sa = pk2(s4.x, s4.y), sb = pk2(s4.z, s4.w);
      const f32x2 dxa = add2(gx2, pk2(nx.x, nx.y)), dxb = add2(gx2, pk2(nx.z, nx.w));
      const f32x2 dya = add2(gy2, pk2(ny.x, ny.y)), dyb = add2(gy2, pk2(ny.z, ny.w));
      const f32x2 sxa = mul2(sa, dxa), sxb = mul2(sb, dxb), sya = mul2(sa, dya), syb = mul2(sb, dyb);
      aMx = add2(add2(aMx, sxa), sxb);
      aMy = add2(add2(aMy, sya), syb);
      if (GEOM) {
        aM0 = add2(add2(aM0, sa), sb);
        aMxx = fma2(sxb, dxb, fma2(sxa, dxa, aMxx));
        aMxy = fma2(sxb, dyb, fma2(sxa, dya, aMxy));
        aMyy = fma2(syb, dyb, fma2(sya, dya, aMyy));
      }
      const f32x2 wa = pk2(w4.x, w4.y), wb = pk2(w4.z, w4.w);
#pragma unroll
      for (int ch = 0; ch < 3; ch++) {
        if (DCOL) {
          const float4 p = lds128(pb_pc + (uint32_t)(ch * ROW * 4) + o);
          aC[ch] = fma2(wb, pk2(p.z, p.w), fma2(wa, pk2(p.x, p.y), aC[ch]));
        }
        if (FEAT) {
          const float4 p = lds128(pb_pc + (uint32_t)((3 + ch) * ROW * 4) + o);
          aF[ch] = fma2(wb, pk2(p.z, p.w), fma2(wa, pk2(p.x, p.y), aF[ch]));
        }
      }
    }
    // the two pixels of each packed accumulator, then the four quarters of the column: two value-halving rounds,
    // after which lane (c, q) holds three finished sums — q = 0: Mx, My, M0;  1: Mxx, Mxy, Myy;  2: colour;  3: feature
    float v[12];
    auto fold = [](f32x2 a) { float lo, hi; upk2(a, lo, hi); return lo + hi; };
    v[0] = fold(aMx); v[1] = fold(aMy); v[2] = fold(aM0);
    v[3] = fold(aMxx); v[4] = fold(aMxy); v[5] = fold(aMyy);
#pragma unroll
    for (int ch = 0; ch < 3; ch++) { v[6 + ch] = fold(aC[ch]); v[9 + ch] = fold(aF[ch]); }
    {
      const bool upper = (lane & 16) != 0;
#pragma unroll
      for (int k = 0; k < 6; k++) {
        const float send = upper ? v[k] : v[k + 6];
        const float keep = upper ? v[k + 6] : v[k];
        v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
      }
    }
    {
      const bool upper = (lane & 8) != 0;
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const float send = upper ? v[k] : v[k + 3];
        const float keep = upper ? v[k + 3] : v[k];
        v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
      }
    }
    if (acc != nullptr) {
      // Packed accumulation: the instance's four lanes each add their three sums (and a zero) to one 16-byte quarter
      // of the Gaussian's 64-byte record with ONE vector RED — 1 instruction and 16 sectors per flush instead of 12
      // instructions and 96 sectors.  ncu (profiles/r2o): the L1 data pipe was 76 % busy, a quarter of its wavefronts
      // the scalar REDs; timing-only experiments (GPU call r2p): no REDs at all 740 us, vector REDs 740 us, scalar
      // REDs 799 us.  The per-Gaussian kernel (or unpack_acc_kernel) reads the record back.
      if (pb_c < n_used) {
        float a0 = v[0], a1 = v[1], a2 = v[2];
        bool live = pb_q == 0 || (pb_q == 1 && GEOM) || (pb_q == 2 && DCOL) || (pb_q == 3 && FEAT);
        if (!GEOM && pb_q == 0) {  // nobody applies the coefficients later: finish dL_dmean2D here (see below)
          const float4 cn = lds128(pb_conic);
          a0 = -m.w * (cn.x * v[0] + cn.y * v[1]) * half_W;
          a1 = -m.w * (cn.z * v[1] + cn.y * v[0]) * half_H;
          a2 = 0.f;
        }
        if (live) {
          float* dst = acc + (size_t)__float_as_uint(m.z) * ACC_FLOATS + pb_q * 4;
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(a0), "f"(a1), "f"(a2), "f"(0.f)
                       : "memory");
        }
      }
      return;
    }
    if (pb_c < n_used) {
      const uint32_t gid = __float_as_uint(m.z);
      if (pb_q == 0) {
        if (GEOM) {  // moments: the per-Gaussian kernel applies the coefficients (preprocess_bwd.cu)
          atomicAdd(dL_dmean2D + (size_t)gid * 3, v[0]);
          atomicAdd(dL_dmean2D + (size_t)gid * 3 + 1, v[1]);
          atomicAdd(dL_dopacity + gid, v[2]);
        } else {  // nobody follows: finish dL_dmean2D = -o * (cx Mx + cy My, cz My + cy Mx) * (W/2, H/2) here
          const float4 cn = lds128(pb_conic);
          atomicAdd(dL_dmean2D + (size_t)gid * 3, -m.w * (cn.x * v[0] + cn.y * v[1]) * half_W);
          atomicAdd(dL_dmean2D + (size_t)gid * 3 + 1, -m.w * (cn.z * v[1] + cn.y * v[0]) * half_H);
        }
      } else if (pb_q == 1) {
        if (GEOM) {
          atomicAdd(dL_dconic + (size_t)gid * 4, v[0]);
          atomicAdd(dL_dconic + (size_t)gid * 4 + 1, v[1]);
          atomicAdd(dL_dconic + (size_t)gid * 4 + 3, v[2]);
        }
      } else if (pb_q == 2) {
        if (DCOL) {
#pragma unroll
          for (int ch = 0; ch < 3; ch++) atomicAdd(dL_dcolors + (size_t)gid * 3 + ch, v[ch]);
        }
      } else {
        if (FEAT) {
#pragma unroll
          for (int ch = 0; ch < 3; ch++) atomicAdd(dL_dfeat + (size_t)gid * 3 + ch, v[ch]);
        }
      }
    }
  };

  // ---- phase A: back to front: the batch starting at `top` covers forward indices top-lane, top = eff_len-1
  // downwards (backward.cu:486-498) ----
  const uint32_t smA = smem_u32(&sm.batch.A[0]);
  const uint32_t stW = smem_u32(&sm.W[0][0]) + (uint32_t)lane * 4u;  // + column * ROW * 4 + slot * 36 * 4
  const uint32_t stS = smem_u32(&sm.S[0][0]) + (uint32_t)lane * 4u;
  const uint32_t stMeta = smem_u32(&sm.meta[0]), stConic = smem_u32(&sm.conic[0]);
  int ncol = 0;
  Fetched nxt = fetch_instance(list, rec, range.x, eff_len - 1 - lane);
#if LSR_BWD2_PIPE
  // Software-pipelined walk (LSR_BWD2_PIPE == 1).  The one-instance-at-a-time loop is a single dependent chain per warp (position -> power
  // -> exp -> alpha -> ballot -> reciprocal -> T -> recurrences -> stores; ncu r2c: 23 % of the stall samples are
  // fixed-latency waits, 19 % MUFU / shared-memory scoreboard, issue slots only 70 % used), and its data-dependent
  // branches keep the compiler from overlapping neighbouring iterations.  Here every iteration runs the FRONT half of
  // instance j+1 (geometry, exp, alpha, contribution test: independent of everything before it) next to the BACK half
  // of instance j (the sequential part: T, colour-behind, w and s), both branch-free — lanes that do not contribute
  // compute and discard — so the two chains sit in one basic block and interleave.
  struct Front {
    float G[BLEND_SLOTS], alpha[BLEND_SLOTS];
    bool contrib[BLEND_SLOTS];
    bool any;
  };
  auto front = [&](int j) {
    Front f;
    const uint32_t ad = smA + (uint32_t)j * (uint32_t)sizeof(float4);
    const float4 a = lds128(ad);
    const float4 bb = lds128(ad + SM_B);  // cz, thr_m, opacity, position
    const uint32_t fwd = __float_as_uint(bb.w);
    const float dx = __fadd_rn(a.x, -pxf);
    const float qx = __fmul_rn(a.z, dx), hx = __fmul_rn(a.w, dx);
    bool any = false;
#pragma unroll
    for (int s = 0; s < BLEND_SLOTS; s++) {
      const float dy = __fadd_rn(a.y, -pyf[s]);
      const float q = __fmaf_rn(dx, qx, __fmul_rn(__fmul_rn(bb.x, dy), dy));
      const float power = __fmaf_rn(q, -0.5f, -__fmul_rn(hx, dy));
      // entries at or beyond n_contrib were never blended (backward.cu:504-506); power > 0 and the alpha threshold as
      // in the forward
      const bool maybe = (fwd < lastc[s]) && !(power > 0.0f);
      f.G[s] = expf(power);
      f.alpha[s] = fminf(0.99f, __fmul_rn(bb.z, f.G[s]));
      f.contrib[s] = maybe && !(f.alpha[s] < kAlphaMin);
      any |= f.contrib[s];
    }
    f.any = any;  // this lane's view; the warp-wide vote follows the back half (so that both halves share a block)
    return f;
  };
  auto back = [&](int j, const Front& f) {
    const uint32_t ad = smA + (uint32_t)j * (uint32_t)sizeof(float4);
    const float4 c = lds128(ad + SM_C);  // r, g, b, f0
    const float4 d = lds128(ad + SM_D);  // f1, f2, id
    const uint32_t colofs = (uint32_t)(ncol * ROW * 4);
#pragma unroll
    for (int s = 0; s < BLEND_SLOTS; s++) {
      const bool ct = f.contrib[s];
      const float one_m_alpha = __fadd_rn(1.f, -f.alpha[s]);
#if LSR_BWD_IEEE_DIV
      const float rcp = __fdiv_rn(1.f, one_m_alpha);
      const float Tn = __fdiv_rn(T[s], one_m_alpha);  // the reference's T = T / (1 - alpha) (backward.cu:521)
#else
      const float rcp = rcp_ftz(one_m_alpha);  // 1 - alpha is in [0.01, 1]
      const float Tn = T[s] * rcp;
#endif
      float cd = 0.f;  // colour . dL_dpixel over all blended channels
      if (DCOL) cd = c.x * dpix[s][0] + c.y * dpix[s][1] + c.z * dpix[s][2];
      if (FEAT) cd += c.w * dpixF[s][0] + d.x * dpixF[s][1] + d.y * dpixF[s][2];
      // "colour behind" recurrence on the dot product (backward.cu:527-538, per channel there)
      const float accn = alpha_last[s] * cd_last[s] + (1.f - alpha_last[s]) * accA[s];
      float dL_dalpha = (cd - accn) * Tn;
      if (DCOL) dL_dalpha += bgT[s] * rcp;  // background term (backward.cu:545-550)
      const float wv = ct ? f.alpha[s] * Tn : 0.f;  // d(channel)/d(colour)
      const float sv = ct ? f.G[s] * dL_dalpha : 0.f;
      T[s] = ct ? Tn : T[s];
      accA[s] = ct ? accn : accA[s];
      cd_last[s] = ct ? cd : cd_last[s];
      alpha_last[s] = ct ? f.alpha[s] : alpha_last[s];
      sts32(stW + colofs + (uint32_t)(s * 36 * 4), wv);
      sts32(stS + colofs + (uint32_t)(s * 36 * 4), sv);
    }
    if (lane == 0) {
      const float4 a = lds128(ad);
      const float4 bb = lds128(ad + SM_B);
      asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(stMeta + (uint32_t)ncol * 16u), "f"(a.x), "f"(a.y),
                   "f"(d.z), "f"(bb.z)
                   : "memory");
      if (!GEOM)
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(stConic + (uint32_t)ncol * 16u), "f"(a.z),
                     "f"(a.w), "f"(bb.x), "f"(0.f)
                     : "memory");
    }
  };
#if LSR_BWD2_PIPE == 2
  for (int top = eff_len - 1; top >= 0; top -= 32) {
    __syncwarp();  // everyone has left the previous batch
    const int nb = commit_batch(sm.batch, nxt, X0, Y0);
    nxt = fetch_instance(list, rec, range.x, top - 32 - lane);
    for (int j = 0; j < nb; j++) {
      const uint32_t ad = smA + (uint32_t)j * (uint32_t)sizeof(float4);
      const float4 a = lds128(ad);
      const float4 bb = lds128(ad + SM_B);  // cz, thr_m, opacity, position
      const uint32_t fwd = __float_as_uint(bb.w);
      const float dx = __fadd_rn(a.x, -pxf);
      const float qx = __fmul_rn(a.z, dx), hx = __fmul_rn(a.w, dx);
      float power[BLEND_SLOTS];
      bool maybe[BLEND_SLOTS], any_maybe = false;
#pragma unroll
      for (int s = 0; s < BLEND_SLOTS; s++) {
        const float dy = __fadd_rn(a.y, -pyf[s]);
        const float q = __fmaf_rn(dx, qx, __fmul_rn(__fmul_rn(bb.x, dy), dy));
        power[s] = __fmaf_rn(q, -0.5f, -__fmul_rn(hx, dy));
        maybe[s] = (fwd < lastc[s]) && !(power[s] > 0.0f) && !(power[s] < bb.y);
        any_maybe |= maybe[s];
      }
      if (!__any_sync(0xffffffffu, any_maybe)) continue;
      Front f;
      bool any = false;
#pragma unroll
      for (int s = 0; s < BLEND_SLOTS; s++) {
        f.G[s] = expf(power[s]);
        f.alpha[s] = fminf(0.99f, __fmul_rn(bb.z, f.G[s]));
        f.contrib[s] = maybe[s] && !(f.alpha[s] < kAlphaMin);
        any |= f.contrib[s];
      }
      if (!__any_sync(0xffffffffu, any)) continue;
      back(j, f);
      if (++ncol == NCOL) {
        __syncwarp();  // the panel is complete
        flush(NCOL);
        ncol = 0;
        __syncwarp();  // ... and free again
      }
    }
  }
#else
  for (int top = eff_len - 1; top >= 0; top -= 32) {
    __syncwarp();  // everyone has left the previous batch
    const int nb = commit_batch(sm.batch, nxt, X0, Y0);
    nxt = fetch_instance(list, rec, range.x, top - 32 - lane);
    if (nb == 0) continue;
    Front cur = front(0);
    cur.any = __any_sync(0xffffffffu, cur.any);
    for (int j = 0; j < nb; j++) {
      // (the last iteration evaluates the front of its own instance once more instead of branching around it)
      Front nxtf = front(min(j + 1, nb - 1));
      back(j, cur);  // a column nobody contributes to is written but not kept
      nxtf.any = __any_sync(0xffffffffu, nxtf.any);
      if (cur.any && ++ncol == NCOL) {
        __syncwarp();  // the panel is complete
        flush(NCOL);
        ncol = 0;
        __syncwarp();  // ... and free again
      }
      cur = nxtf;
    }
  }
#endif
#else
  for (int top = eff_len - 1; top >= 0; top -= 32) {
    __syncwarp();  // everyone has left the previous batch
    const int nb = commit_batch(sm.batch, nxt, X0, Y0);
    nxt = fetch_instance(list, rec, range.x, top - 32 - lane);
#if LSR_BWD2_PREFETCH
    float4 a_nx = lds128(smA), bb_nx = lds128(smA + SM_B);  // record of instance 0 (garbage if nb == 0: unused)
#endif
    for (int j = 0; j < nb; j++) {
      const uint32_t ad = smA + (uint32_t)j * (uint32_t)sizeof(float4);
#if LSR_BWD2_PREFETCH
      const float4 a = a_nx, bb = bb_nx;
      a_nx = lds128(ad + (uint32_t)sizeof(float4));  // the next instance's record: its latency hides behind this one
      bb_nx = lds128(ad + (uint32_t)sizeof(float4) + SM_B);
#else
      const float4 a = lds128(ad);
      const float4 bb = lds128(ad + SM_B);  // cz, thr_m, opacity, position
#endif
#if LSR_BWD2_EARLY_CD
      const float4 c = lds128(ad + SM_C);  // r, g, b, f0: requested with the geometry, not after the votes
      const float4 d = lds128(ad + SM_D);  // f1, f2, id
#endif
      const uint32_t fwd = __float_as_uint(bb.w);
      const float dx = __fadd_rn(a.x, -pxf);
      const float qx = __fmul_rn(a.z, dx), hx = __fmul_rn(a.w, dx);
      float power[BLEND_SLOTS];
      bool maybe[BLEND_SLOTS];
#pragma unroll
      for (int s = 0; s < BLEND_SLOTS; s++) {
        const float dy = __fadd_rn(a.y, -pyf[s]);
        const float q = __fmaf_rn(dx, qx, __fmul_rn(__fmul_rn(bb.x, dy), dy));
        power[s] = __fmaf_rn(q, -0.5f, -__fmul_rn(hx, dy));
        // entries at or beyond n_contrib were never blended (backward.cu:504-506); power > 0 and the alpha threshold
        // as in the forward.  The pre-test against thr_m avoids expf for pairs that certainly fail alpha >= 1/255.
        maybe[s] = (fwd < lastc[s]) && !(power[s] > 0.0f) && !(power[s] < bb.y);
      }
#if !LSR_BWD2_ONE_VOTE
      bool any_maybe = false;
#pragma unroll
      for (int s = 0; s < BLEND_SLOTS; s++) any_maybe |= maybe[s];
      if (!__any_sync(0xffffffffu, any_maybe)) continue;
#endif
      float G[BLEND_SLOTS], alpha[BLEND_SLOTS];
      bool contrib[BLEND_SLOTS];
#if LSR_BWD2_ONE_VOTE
      bool mine = false;  // one warp-wide vote per instance instead of three: every vote is a stop in the dependent chain
#pragma unroll
      for (int s = 0; s < BLEND_SLOTS; s++) {
        G[s] = bwd_exp(power[s]);
        alpha[s] = fminf(0.99f, __fmul_rn(bb.z, G[s]));
      }
#if LSR_BWD2_FAST_EXP == 2
      if (fabsf(alpha[0] - kAlphaMin) < kAlphaBand || fabsf(alpha[1] - kAlphaMin) < kAlphaBand) {
#pragma unroll
        for (int s = 0; s < BLEND_SLOTS; s++) {
          G[s] = expf(power[s]);
          alpha[s] = fminf(0.99f, __fmul_rn(bb.z, G[s]));
        }
      }
#endif
#pragma unroll
      for (int s = 0; s < BLEND_SLOTS; s++) {
        contrib[s] = maybe[s] && !(alpha[s] < kAlphaMin);
        mine |= contrib[s];
      }
#if LSR_BWD2_ONE_VOTE == 1
      if (!__any_sync(0xffffffffu, mine)) continue;
#endif  // (== 2: no vote at all; an instance nobody contributes to still takes a column of zeros)
#else
      uint32_t call = 0u;
#pragma unroll
      for (int s = 0; s < BLEND_SLOTS; s++) {
        G[s] = expf(power[s]);
        alpha[s] = fminf(0.99f, __fmul_rn(bb.z, G[s]));
        contrib[s] = maybe[s] && !(alpha[s] < kAlphaMin);
        call |= __ballot_sync(0xffffffffu, contrib[s]);
      }
      if (call == 0u) continue;
#endif

#if !LSR_BWD2_EARLY_CD
      const float4 c = lds128(ad + SM_C);  // r, g, b, f0
      const float4 d = lds128(ad + SM_D);  // f1, f2, id
#endif
      const uint32_t colofs = (uint32_t)(ncol * ROW * 4);
#pragma unroll
      for (int s = 0; s < BLEND_SLOTS; s++) {
        float wv = 0.f, sv = 0.f;
        if (contrib[s]) {
          const float one_m_alpha = __fadd_rn(1.f, -alpha[s]);
#if LSR_BWD_IEEE_DIV
          const float rcp = __fdiv_rn(1.f, one_m_alpha);
          T[s] = __fdiv_rn(T[s], one_m_alpha);  // the reference's T = T / (1 - alpha) (backward.cu:521)
#else
          const float rcp = rcp_ftz(one_m_alpha);  // 1 - alpha is in [0.01, 1]
          T[s] *= rcp;
#endif
          wv = alpha[s] * T[s];  // d(channel)/d(colour)
          float cd = 0.f;        // colour . dL_dpixel over all blended channels
          if (DCOL) cd = c.x * dpix[s][0] + c.y * dpix[s][1] + c.z * dpix[s][2];
          if (FEAT) cd += c.w * dpixF[s][0] + d.x * dpixF[s][1] + d.y * dpixF[s][2];
          // "colour behind" recurrence on the dot product (backward.cu:527-538, per channel there)
          accA[s] = alpha_last[s] * cd_last[s] + (1.f - alpha_last[s]) * accA[s];
          cd_last[s] = cd;
          alpha_last[s] = alpha[s];
          float dL_dalpha = (cd - accA[s]) * T[s];
          if (DCOL) dL_dalpha += bgT[s] * rcp;  // background term (backward.cu:545-550)
          sv = G[s] * dL_dalpha;
        }
        sts32(stW + colofs + (uint32_t)(s * 36 * 4), wv);
        sts32(stS + colofs + (uint32_t)(s * 36 * 4), sv);
      }
      if (lane == 0) {
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(stMeta + (uint32_t)ncol * 16u), "f"(a.x), "f"(a.y),
                     "f"(d.z), "f"(bb.z)
                     : "memory");
        if (!GEOM)
          asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(stConic + (uint32_t)ncol * 16u), "f"(a.z),
                       "f"(a.w), "f"(bb.x), "f"(0.f)
                       : "memory");
      }
      if (++ncol == NCOL) {
        __syncwarp();  // the panel is complete
        flush(NCOL);
        ncol = 0;
        __syncwarp();  // ... and free again
      }
    }
  }
#endif
  if (ncol > 0) {
    __syncwarp();
    flush(ncol);
  }
}

template <bool FEAT, bool DCOL, bool GEOM>
void launch_one(unsigned grid, cudaStream_t s, int W, int H, int tiles_x, const uint2* ranges, const uint32_t* order,
                const uint32_t* list, const GeomRec* rec, const float* bg, const float* final_T, const uint32_t* n_contrib,
                const float* dL_dcolor, const float* dL_dfeature, float* dL_dmean2D, float* dL_dconic,
                float* dL_dopacity, float* dL_dcolors, float* dL_dfeat, float* acc) {
  render_bwd2_kernel<FEAT, DCOL, GEOM><<<grid, 32, 0, s>>>(W, H, tiles_x, ranges, order, list, rec, bg, final_T,
                                                         n_contrib, dL_dcolor, dL_dfeature, dL_dmean2D, dL_dconic,
                                                         dL_dopacity, dL_dcolors, dL_dfeat, acc);
}

}  // namespace

void launch_render_bwd2(int W, int H, const uint2* ranges, const uint32_t* order, const uint32_t* list,
                        const GeomRec* rec, const float* bg, bool include_feature, const float* final_T,
                        const uint32_t* n_contrib, const float* dL_dcolor, const float* dL_dfeature, uint32_t need_mask,
                        float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolors, float* dL_dfeat,
                        float* acc, cudaStream_t s) {
  const int tiles_x = (W + TILE_X - 1) / TILE_X, tiles_y = (H + TILE_Y - 1) / TILE_Y;
  if (tiles_x == 0 || tiles_y == 0) return;
  const unsigned grid = (unsigned)tiles_x * (unsigned)tiles_y * QUADS;
  const bool feat = include_feature && dL_dfeature != nullptr;
  const bool dcol = dL_dcolor != nullptr;
  const bool geom = (need_mask & LSR_NEED_GEOMETRY) != 0;
  if (!feat && !dcol) return;  // no upstream gradient at all: everything stays zero
#define LSR_BWD(F, D, G)                                                                                      \
  launch_one<F, D, G>(grid, s, W, H, tiles_x, ranges, order, list, rec, bg, final_T, n_contrib, dL_dcolor,   \
                      dL_dfeature, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolors, dL_dfeat, acc)
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
#endif  // LSR_MOMENTS
