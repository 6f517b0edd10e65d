// decoder.cu — eval-side decode of rendered language features: the autoencoder's decoder MLP
// (autoencoder/model.py:18-25,42-46: Linear, (ReLU, Linear)*, x / ||x||; 3 -> 16 -> 32 -> 64 -> 128 -> 256 -> 256 -> 512
// with the defaults of eval/evaluate_iou_loc.py) followed by the CLIP relevancy of every pixel against the phrase
// embeddings (eval/openclip_encoder.py:42-56, 96-111).  This is the one dense contraction next to the rasterizer.
// The reference runs seven cuBLAS SGEMMs with [N,16..512] fp32 activations round-tripping HBM (about 10 KB per pixel)
// and then an [N,512] x [512,n] mm per level; here ONE kernel keeps a tile of pixels on chip from the 3-d input to
// the relevancy value: 12 B read and 4*n_pos B written per pixel (plus 2 KB if the restored feature is wanted).
//
// Kernels behind lsr_decode:
//   * decode_tc_kernel<F16> (LSR_DECODE_TF32 / LSR_DECODE_F16): 128 pixels per CTA = the 128 TMEM lanes.  Every layer
//     is a tcgen05.mma chain (kind::tf32, K = 8 per instruction, or kind::f16, K = 16; M = 128, N = layer width) issued
//     by one elected lane of warp 0; the A operand (activations, in the canonical no-swizzle K-major core-matrix layout:
//     [K chunks of 16 B][128 rows]) stays in shared memory, the B operand (weights, pre-packed once by lsr_decoder_pack
//     into the same layout) streams from L2 through a ring of 16 KB slices filled by cp.async.bulk (TMA 1-D) on
//     mbarriers, the accumulator lives in TMEM (256 columns; a 512-wide last layer is two passes).  Epilogue per layer:
//     tcgen05.ld, + bias, ReLU, back into the A buffer; last layer: norm straight out of TMEM; the phrase similarities
//     are one more N = 16 pass against P W (fold_phrases_kernel), then the softmax pair on the CUDA cores.  TF32: fp32 activations (128 KB), one CTA per SM, operands rounded to the
//     nearest TF32 value (the tensor core then truncates exactly).  F16: activations and weights rounded to nearest fp16 (same 10 mantissa bits,
//     values clamped to +-65504), 64 KB of activations, two CTAs per SM so one CTA's epilogue overlaps the other's
//     MMAs, half the weight bytes per tile and twice the MMA rate.  fp32 accumulation, bias, norm and relevancy in both.
//   * decode_fp32_kernel (LSR_DECODE_FP32): the same fusion on the FP32 pipe (32 pixels per CTA, activations ping-pong
//     in shared memory), IEEE fp32 like the reference's SGEMM — the exact-arithmetic variant.
#include <cuda_fp16.h>

#include <cmath>
#include <cstdlib>

#include "common.cuh"

namespace lsr {

namespace {

constexpr int DEC_MAXL = LSR_DECODER_MAX_LAYERS;
constexpr int DEC_MAXW = 512;
constexpr int DEC_MAXPH = LSR_DECODER_MAX_PHRASES;

// ------------------------------------------------------------------------------------------------------------------
// relevancy of one pixel from its similarities (softmax over the pair (10 pos, 10 neg_j), the j with the lowest
// positive probability = the largest negative similarity)
__device__ __forceinline__ float pair_softmax_pos(float pos, float max_neg) {
  const float a = 10.0f * pos, b = 10.0f * max_neg;
  const float m = fmaxf(a, b);
  const float e0 = expf(a - m), e1 = expf(b - m);
  return e0 / (e0 + e1);
}

// ------------------------------------------------------------------------------------------------------------------
// FP32 variant
constexpr int F_TP = 32;        // pixels per CTA
constexpr int F_THREADS = 256;
constexpr int F_LD = DEC_MAXW;  // row stride of an activation buffer

struct DecArgs {
  int n_layers;
  int dims[DEC_MAXL + 1];
  const float* weight[DEC_MAXL];
  const float* bias[DEC_MAXL];
};

__global__ void __launch_bounds__(F_THREADS)
decode_fp32_kernel(const DecArgs a, const float* __restrict__ in, int64_t n, float* __restrict__ out,
                   const float* __restrict__ phrases, int n_pos, int n_neg, float* __restrict__ relev) {
  extern __shared__ __align__(16) float fsm[];
  float* cur = fsm;
  float* nxt = fsm + F_TP * F_LD;
  const int tid = threadIdx.x;
  const int64_t ntiles = (n + F_TP - 1) / F_TP;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * F_TP;
    const int d0 = a.dims[0], d0p = (d0 + 3) & ~3;
    for (int i = tid; i < F_TP * d0p; i += F_THREADS) {
      const int p = i / d0p, k = i - p * d0p;
      cur[p * F_LD + k] = (k < d0 && base + p < n) ? __ldg(in + (base + p) * d0 + k) : 0.0f;
    }
    __syncthreads();
    const int pg = tid >> 6, nl = tid & 63;  // 4 groups of 8 pixels x 64 output lanes
    for (int l = 0; l < a.n_layers; l++) {
      const int K = a.dims[l], N = a.dims[l + 1];
      const float* __restrict__ W = a.weight[l];
      const float* __restrict__ B = a.bias[l];
      const bool relu = l + 1 < a.n_layers;
      const bool vec = (K & 3) == 0 && (((uintptr_t)W) & 15) == 0;
      const float* act = cur + pg * 8 * F_LD;
      for (int o = nl; o < N; o += 64) {
        float acc[8];
        const float b = __ldg(B + o);
#pragma unroll
        for (int p = 0; p < 8; p++) acc[p] = b;
        const float* w = W + (size_t)o * K;
        for (int k = 0; k < K; k += 4) {
          float4 wv;
          if (vec) {
            wv = __ldg(reinterpret_cast<const float4*>(w + k));
          } else {
            wv.x = __ldg(w + k);
            wv.y = k + 1 < K ? __ldg(w + k + 1) : 0.0f;
            wv.z = k + 2 < K ? __ldg(w + k + 2) : 0.0f;
            wv.w = k + 3 < K ? __ldg(w + k + 3) : 0.0f;
          }
#pragma unroll
          for (int p = 0; p < 8; p++) {
            const float4 x = *reinterpret_cast<const float4*>(act + p * F_LD + k);
            acc[p] = fmaf(x.x, wv.x, acc[p]);
            acc[p] = fmaf(x.y, wv.y, acc[p]);
            acc[p] = fmaf(x.z, wv.z, acc[p]);
            acc[p] = fmaf(x.w, wv.w, acc[p]);
          }
        }
#pragma unroll
        for (int p = 0; p < 8; p++) nxt[(pg * 8 + p) * F_LD + o] = relu ? fmaxf(acc[p], 0.0f) : acc[p];
      }
      if (N & 3) {  // zero the pad the next layer's float4 reads touch
        for (int i = tid; i < F_TP * 4; i += F_THREADS) {
          const int p = i >> 2, k = N + (i & 3);
          if (k < ((N + 3) & ~3)) nxt[p * F_LD + k] = 0.0f;
        }
      }
      __syncthreads();
      float* t = cur; cur = nxt; nxt = t;
    }
    // normalisation, output, relevancy: one warp per 4 pixels
    const int NL = a.dims[a.n_layers];
    const int warp = tid >> 5, lane = tid & 31;
    const int n_ph = phrases ? n_pos + n_neg : 0;
    for (int pp = 0; pp < 4; pp++) {
      const int p = warp * 4 + pp;
      const int64_t pix = base + p;
      if (pix >= n) break;
      const float* x = cur + p * F_LD;
      float ss = 0.0f;
      for (int k = lane; k < NL; k += 32) ss = fmaf(x[k], x[k], ss);
#pragma unroll
      for (int s = 16; s > 0; s >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, s);
      const float norm = sqrtf(ss);
      if (out)
        for (int k = lane; k < NL; k += 32) out[pix * NL + k] = x[k] / norm;
      if (n_ph) {
        float max_neg = -INFINITY, mine = 0.0f;  // lane j keeps the similarity of positive j
        for (int j = 0; j < n_ph; j++) {
          const float* ph = phrases + (size_t)j * NL;
          float d = 0.0f;
          for (int k = lane; k < NL; k += 32) d = fmaf(x[k] / norm, __ldg(ph + k), d);
#pragma unroll
          for (int s = 16; s > 0; s >>= 1) d += __shfl_xor_sync(0xffffffffu, d, s);
          if (j < n_pos) { if (lane == j) mine = d; }
          else max_neg = fmaxf(max_neg, d);
        }
        if (lane < n_pos) relev[(size_t)lane * n + pix] = pair_softmax_pos(mine, max_neg);
      }
    }
    __syncthreads();
  }
}
// ------------------------------------------------------------------------------------------------------------------
// Tensor-core variants.  A "pass" is one accumulator's worth of a layer: rows [n0, n0 + Np) of its weight matrix,
// Np <= 256 TMEM columns.  Every layer but the last is one pass (widths <= 256); a last layer wider than 256 is two
// passes of half the width, run one after the other (norm and phrase dot products accumulate across them).
// Packed weights of a pass: [Kp/E][Np][E] elements, E = elements per 16 bytes (4 tf32 / 8 f16), Kp = K rounded up to
// 2E and zero padded: the canonical no-swizzle K-major UMMA layout (8-row x 16-byte core matrices, 128 B apart along
// N, Np*16 B apart along K).
constexpr int TC_M = 128;
#ifndef TC_TF32_STAGE_BYTES
#define TC_TF32_STAGE_BYTES 32768  // measured: 16 KB slices 3.30 ms, 32 KB slices 2.61 ms (per-slice latency amortised)
#endif
constexpr int TC_MAXK = 256;
constexpr int TC_MAX_SLICES = 96;
constexpr int TC_MAX_PASSES = DEC_MAXL + 4;
enum { PASS_HIDDEN = 0, PASS_SIM = 1, PASS_NORM = 2, PASS_STORE = 3 };
constexpr int TC_TMEM_COLS = 256;
#ifndef TC_F16_STAGGER
#define TC_F16_STAGGER 12000  // measured (f16, 1440x1080): 0 -> 1.083 ms, 6000 -> 1.070, 12000 -> 1.003, 20000 -> 1.027
#endif

// Operand modes of the tensor-core kernel.  TC_BX2 ("split bf16"): every fp32 operand x is carried as two bf16 numbers,
// hi = bf16(x) and lo = bf16(x - hi) (16 mantissa bits together), and every product as three bf16 MMAs with fp32
// accumulation, hi*hi + lo*hi + hi*lo (the lo*lo term is 2^-18 relative) — fp32-class results (the reference runs fp32
// SGEMMs) at tensor-core speed: the same operand bytes as TF32, 1.5x its MMA instructions.
enum { TC_TF32 = 0, TC_F16 = 1, TC_BX2 = 2 };
template <int MODE>
struct TcCfg {
  static constexpr bool F16 = MODE == TC_F16, BX2 = MODE == TC_BX2;
  static constexpr int E = MODE == TC_TF32 ? 4 : 8;           // elements per 16-byte K chunk
  static constexpr int PARTS = BX2 ? 2 : 1;                   // operand parts (hi, lo)
  static constexpr int A_PART_BYTES = TC_M * TC_MAXK * (MODE == TC_TF32 ? 4 : 2);
  static constexpr int A_BYTES = A_PART_BYTES * PARTS;
  static constexpr int STAGE_BYTES = F16 ? 16384 : TC_TF32_STAGE_BYTES;
  static constexpr int STAGES = 3;                            // f16: 64 + 48 = 112 KB per CTA -> two CTAs per SM
  static constexpr int BAR_OFF = A_BYTES + STAGES * STAGE_BYTES;
  static constexpr int SMEM = BAR_OFF + 128;
  static constexpr int CTAS_PER_SM = F16 ? 2 : 1;
};

// fp32 -> (hi, lo) bf16 pair, round to nearest even; bf16 keeps fp32's exponent range, so nothing is clamped
__device__ __forceinline__ uint32_t bf16_rn_bits(float x) {
  const uint32_t u = __float_as_uint(x);
  return (u + 0x7FFFu + ((u >> 16) & 1u)) >> 16;
}
__device__ __forceinline__ void bf16_split(float x, uint32_t& hi, uint32_t& lo) {
  hi = bf16_rn_bits(x);
  lo = bf16_rn_bits(x - __uint_as_float(hi << 16));
}
// eight floats -> one 16-byte chunk of hi parts and one of lo parts
__device__ __forceinline__ void bf16_split8(const float* v, uint4& hi, uint4& lo) {
  uint32_t h[8], l[8];
#pragma unroll
  for (int j = 0; j < 8; j++) bf16_split(v[j], h[j], l[j]);
  hi = make_uint4(h[0] | (h[1] << 16), h[2] | (h[3] << 16), h[4] | (h[5] << 16), h[6] | (h[7] << 16));
  lo = make_uint4(l[0] | (l[1] << 16), l[2] | (l[3] << 16), l[4] | (l[5] << 16), l[6] | (l[7] << 16));
}

// fp32 -> nearest TF32 (10 mantissa bits), so the tensor core's truncation of the operands is exact
__device__ __forceinline__ uint32_t tf32_rn(float x) { return (__float_as_uint(x) + 0x1000u) & 0xFFFFE000u; }

// mode TC_BX2: the hi and lo chunks of K chunk kc sit next to each other, [Kp/8][2][Np] 16-byte chunks, so that a K
// slice (hi and lo) is one contiguous run of bytes for the bulk copy
__global__ void pack_weights_kernel(const float* __restrict__ W, int K, int Np, int Kp, int mode, uint4* __restrict__ dst) {
  const int E = mode == TC_TF32 ? 4 : 8;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (Kp / E) * Np) return;
  const int kc = i / Np, o = i - kc * Np;
  float v[8];
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const int k = kc * E + j;
    v[j] = (j < E && k < K) ? __ldg(W + (size_t)o * K + k) : 0.0f;
  }
  if (mode == TC_BX2) {
    uint4 hi, lo;
    bf16_split8(v, hi, lo);
    dst[(size_t)(2 * kc) * Np + o] = hi;
    dst[(size_t)(2 * kc + 1) * Np + o] = lo;
    return;
  }
  uint4 r;
  if (mode == TC_F16) {
    __half2 h0 = __floats2half2_rn(v[0], v[1]), h1 = __floats2half2_rn(v[2], v[3]);
    __half2 h2 = __floats2half2_rn(v[4], v[5]), h3 = __floats2half2_rn(v[6], v[7]);
    r = make_uint4(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1),
                   *reinterpret_cast<uint32_t*>(&h2), *reinterpret_cast<uint32_t*>(&h3));
  } else {
    r = make_uint4(tf32_rn(v[0]), tf32_rn(v[1]), tf32_rn(v[2]), tf32_rn(v[3]));
  }
  dst[i] = r;
}

// Phrase similarities folded into the last layer: sim_j = x . p_j = h . (P W)_j + P_j . b for the last Linear (W, b)
// and its input h, so the similarities of a tile are one more N = 16 MMA pass against PW = P W  [16, K] instead of
// n_ph x NL FMAs per pixel on the FP32 pipe.  Writes PW in the packed operand layout [Kp/E][16][E] (rows past n_ph are
// zeros) and pb[16] = P b.
__global__ void fold_phrases_kernel(const float* __restrict__ P, int n_ph, const float* __restrict__ W,
                                    const float* __restrict__ b, int NL, int K, int Kp, int mode,
                                    uint4* __restrict__ dst, float* __restrict__ pb) {
  const int E = mode == TC_TF32 ? 4 : 8;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int items = (Kp / E) * 16;
  if (i < items) {
    const int kc = i >> 4, j = i & 15;
    float acc[8] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    if (j < n_ph) {
      const float* pj = P + (size_t)j * NL;
      for (int o = 0; o < NL; o++) {
        const float pv = __ldg(pj + o);
        const float* w = W + (size_t)o * K + kc * E;
#pragma unroll
        for (int e = 0; e < 8; e++)
          if (e < E && kc * E + e < K) acc[e] = fmaf(pv, __ldg(w + e), acc[e]);
      }
    }
    uint4 r;
    if (mode == TC_BX2) {  // [Kp/8][2][16] chunks, as pack_weights_kernel lays a pass out
      uint4 hi, lo;
      bf16_split8(acc, hi, lo);
      dst[(2 * kc) * 16 + j] = hi;
      dst[(2 * kc + 1) * 16 + j] = lo;
      return;
    }
    if (mode == TC_F16) {
      __half2 h0 = __floats2half2_rn(acc[0], acc[1]), h1 = __floats2half2_rn(acc[2], acc[3]);
      __half2 h2 = __floats2half2_rn(acc[4], acc[5]), h3 = __floats2half2_rn(acc[6], acc[7]);
      r = make_uint4(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1),
                     *reinterpret_cast<uint32_t*>(&h2), *reinterpret_cast<uint32_t*>(&h3));
    } else {
      r = make_uint4(tf32_rn(acc[0]), tf32_rn(acc[1]), tf32_rn(acc[2]), tf32_rn(acc[3]));
    }
    dst[i] = r;
  } else if (i < items + 16) {
    const int j = i - items;
    float acc = 0.0f;
    if (j < n_ph)
      for (int o = 0; o < NL; o++) acc = fmaf(__ldg(P + (size_t)j * NL + o), __ldg(b + o), acc);
    pb[j] = acc;
  }
}

struct TcSlice {
  uint32_t off16;  // offset of the slice in the packed blob, 16-byte units
  uint16_t nch;    // 16-byte K chunks in the slice (even)
  uint8_t c0;      // first chunk index within the pass
  uint8_t pass;
};
struct TcArgs {
  int n_layers;
  int dims[DEC_MAXL + 1];
  const float* bias[DEC_MAXL];
  const void* packed;
  const void* fold;    // PW operand of the similarity pass (pass_layer == n_layers), null without phrases
  const float* fold_b; // P b  [16]
  int n_passes;
  int pass_layer[TC_MAX_PASSES], pass_n0[TC_MAX_PASSES], pass_np[TC_MAX_PASSES];
  int pass_mode[TC_MAX_PASSES];  // PASS_HIDDEN / PASS_SIM / PASS_NORM (last layer: sum of squares) / PASS_STORE
  int last_norm_pass;            // the norm (and the relevancy) is final after this pass
  int stagger;                   // cycles the second CTA of an SM waits before its first tile (see launch_tc)
  int pass_slice_end[TC_MAX_PASSES];  // one past the last slice of each pass
  int n_slices;
  TcSlice sl[TC_MAX_SLICES];
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Bounded spin: a protocol bug traps (CUDA error on the host) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  for (uint32_t spin = 0; !ok; spin++) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (spin > (1u << 26)) __trap();
  }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
template <int MODE>
__device__ __forceinline__ void tc_mma(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                       uint32_t accumulate) {
  if constexpr (MODE != TC_TF32) {  // fp16 and bf16 operands both run as kind::f16 (the formats are in idesc)
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
// K-major, no swizzle: start address, LBO = byte distance of the two 16-byte K chunks of one instruction,
// SBO = byte distance of consecutive 8-row groups, descriptor version 1 (sm_100).
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
// 16 consecutive accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, "
      "[%16];\n\t"
      "tcgen05.wait::ld.sync.aligned;"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
#pragma unroll
  for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
}
// One lane of a converged warp (the same one every time: tcgen05.commit tracks the MMAs of the thread that issued them)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
  __half2 h = __floats2half2_rn(fminf(a, 65504.0f), fminf(b, 65504.0f));
  return *reinterpret_cast<uint32_t*>(&h);
}

template <int MODE>
__global__ void __launch_bounds__(TC_M, TcCfg<MODE>::CTAS_PER_SM)
decode_tc_kernel(const __grid_constant__ TcArgs a, const float* __restrict__ in, int64_t n, float* __restrict__ out,
                 const float* __restrict__ phrases, int n_pos, int n_neg, float* __restrict__ relev,
                 long long* __restrict__ trace) {
  using C = TcCfg<MODE>;
  constexpr int E = C::E;
  constexpr bool F16 = C::F16, BX2 = C::BX2;
  extern __shared__ __align__(1024) uint8_t tsm[];
  uint4* A = reinterpret_cast<uint4*>(tsm);  // [K/E][128 rows] 16-byte chunks (TC_BX2: the hi parts)
  uint4* A_lo = reinterpret_cast<uint4*>(tsm + C::A_PART_BYTES);  // TC_BX2: the lo parts, same layout
  uint8_t* stages = tsm + C::A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(tsm + C::BAR_OFF);  // full[STAGES], empty[STAGES], done
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * C::STAGES + 1);
  const int tid = threadIdx.x, warp = tid >> 5;
  // warp index as a value the compiler knows to be warp-uniform: the producer and the issuer run as whole warps with
  // uniform descriptors (in uniform registers) and elect one lane per instruction; under `if (tid == 0)` every
  // tcgen05.mma was wrapped in an ELECT / R2UR.BROADCAST waterfall (~140 cycles per instruction, measured)
  const int warp_u = __shfl_sync(0xffffffffu, warp, 0);
  const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + C::STAGES), done = smem_u32(bars + 2 * C::STAGES);
  const uint32_t a_base = smem_u32(A), st_base = smem_u32(stages);
  const uint8_t* packed = reinterpret_cast<const uint8_t*>(a.packed);
  const int NL = a.dims[a.n_layers];
  const int n_ph = phrases ? n_pos + n_neg : 0;

  if (tid == 0) {
    for (int i = 0; i < C::STAGES; i++) { mbar_init(full0 + 8 * i, 1); mbar_init(empty0 + 8 * i, 1); }
    mbar_init(done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "n"(TC_TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t t_lane = tmem + ((uint32_t)(warp * 32) << 16);  // this warp's quarter of the 128 lanes

  if (a.stagger > 0 && blockIdx.x >= (gridDim.x + 1) / 2) {
    const long long t0 = clock64();
    while (clock64() - t0 < a.stagger) {}
  }
  uint32_t it_load = 0, it_mma = 0;  // warp 1 / warp 0: slices requested / multiplied so far (ring position, phase)
  uint32_t done_phase = 0;
  const int64_t ntiles = (n + TC_M - 1) / TC_M;
  const uint32_t total_loads = (uint32_t)((ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x) * (uint32_t)a.n_slices;
  int next_load = 0;         // warp 1: table index of the next slice to request (wraps at the end of a tile)
  uint32_t tile_iter = 0;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, tile_iter++) {
    const int64_t pix = tile * TC_M + tid;
    const bool live = pix < n;
    {  // input row -> A, K padded to 2E with zeros
      const int d0 = a.dims[0], kc_n = ((d0 + 2 * E - 1) & ~(2 * E - 1)) / E;
      for (int kc = 0; kc < kc_n; kc++) {
        float v[E];
#pragma unroll
        for (int j = 0; j < E; j++) {
          const int k = kc * E + j;
          v[j] = (live && k < d0) ? __ldg(in + pix * d0 + k) : 0.0f;
        }
        if constexpr (BX2) {
          float v8[8];
#pragma unroll
          for (int j = 0; j < 8; j++) v8[j] = v[j % E];
          uint4 hi, lo;
          bf16_split8(v8, hi, lo);
          A[kc * TC_M + tid] = hi;
          A_lo[kc * TC_M + tid] = lo;
        } else if constexpr (F16) {
          auto cv = [](float x, float y) {
            __half2 h = __floats2half2_rn(fminf(fmaxf(x, -65504.0f), 65504.0f), fminf(fmaxf(y, -65504.0f), 65504.0f));
            return *reinterpret_cast<uint32_t*>(&h);
          };
          A[kc * TC_M + tid] = make_uint4(cv(v[0], v[1]), cv(v[2], v[3]), cv(v[4 % E], v[5 % E]), cv(v[6 % E], v[7 % E]));
        } else {
          A[kc * TC_M + tid] = make_uint4(tf32_rn(v[0]), tf32_rn(v[1]), tf32_rn(v[2]), tf32_rn(v[3]));
        }
      }
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    const bool tr = trace && blockIdx.x == 0 && tid == 0 && tile == (int64_t)gridDim.x;  // CTA 0's second tile
    int ti = 0;
    if (tr) trace[ti++] = clock64();
    float ssp[4] = {0.0f, 0.0f, 0.0f, 0.0f}, sim[16], norm = 1.0f;
#pragma unroll
    for (int j = 0; j < 16; j++) sim[j] = 0.0f;

    for (int p = 0; p < a.n_passes; p++) {
      const int l = a.pass_layer[p], n0 = a.pass_n0[p], Np = a.pass_np[p];
      const int s_begin = p ? a.pass_slice_end[p - 1] : 0, s_end = a.pass_slice_end[p];
      if (warp_u == 1) {
        // producer warp: weight slices of this pass, and up to STAGES - 1 slices of the next ones while this pass's
        // epilogue runs (a stage is reused once the MMAs that read it have completed: empty barrier)
        const uint32_t target = min(total_loads, tile_iter * (uint32_t)a.n_slices + (uint32_t)(s_end + C::STAGES - 1));
        while (it_load < target) {  // runs into the next tile's first slices at the end of this one
          const TcSlice q = a.sl[next_load];
          const uint32_t st = it_load % C::STAGES, use = it_load / C::STAGES;
          mbar_wait(empty0 + 8 * st, (use & 1) ^ 1);
          const uint32_t bytes = (uint32_t)q.nch * (uint32_t)a.pass_np[q.pass] * 16u * (uint32_t)C::PARTS;
          const uint8_t* src = a.pass_layer[q.pass] == a.n_layers ? reinterpret_cast<const uint8_t*>(a.fold) : packed;
          if (elect_one()) {
            mbar_expect_tx(full0 + 8 * st, bytes);
            bulk_g2s(st_base + st * C::STAGE_BYTES, src + (size_t)q.off16 * 16, bytes, full0 + 8 * st);
          }
          __syncwarp();
          it_load++;
          if (++next_load == a.n_slices) next_load = 0;
        }
      } else if (warp_u == 0) {
        // MMA issuer warp
        // operand formats: kind::tf32 -> 2 (TF32); kind::f16 -> 0 (F16) or 1 (BF16)
        const uint32_t fmt = MODE == TC_TF32 ? 2u : (BX2 ? 1u : 0u);
        const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(Np >> 3) << 17) | ((TC_M >> 4) << 24);
        // byte distance of the two 16-byte K chunks of one instruction (TC_BX2: a weight chunk's hi and lo parts sit
        // next to each other, so consecutive K chunks of one part are two rows of Np chunks apart)
        const uint32_t a_lbo = TC_M * 16, b_lbo = (uint32_t)Np * 16 * (uint32_t)C::PARTS, sbo = 128;
        const uint64_t a_step = (uint64_t)((2 * a_lbo) >> 4), b_step = (uint64_t)((2 * b_lbo) >> 4);
        const uint64_t a_lo_ofs = (uint64_t)(C::A_PART_BYTES >> 4), b_lo_ofs = (uint64_t)(((uint32_t)Np * 16) >> 4);
        for (int s = s_begin; s < s_end; s++) {
          const TcSlice q = a.sl[s];
          const uint32_t st = it_mma % C::STAGES, use = it_mma / C::STAGES;
          uint64_t ad = smem_desc(a_base + (uint32_t)q.c0 * a_lbo, a_lbo, sbo);
          uint64_t bd = smem_desc(st_base + st * C::STAGE_BYTES, b_lbo, sbo);
          uint32_t acc = q.c0 > 0 ? 1u : 0u;
          const int steps = q.nch >> 1;
          mbar_wait(full0 + 8 * st, use & 1);
          tc_fence_after();
          if (elect_one()) {
#pragma unroll 4
            for (int j = 0; j < steps; j++) {
              tc_mma<MODE>(tmem, ad, bd, idesc, acc);  // (hi) x (hi)
              acc = 1u;
              if constexpr (BX2) {
                tc_mma<MODE>(tmem, ad + a_lo_ofs, bd, idesc, 1u);  // lo x hi
                tc_mma<MODE>(tmem, ad, bd + b_lo_ofs, idesc, 1u);  // hi x lo
              }
              ad += a_step;
              bd += b_step;
            }
            tc_commit(empty0 + 8 * st);  // the slice's stage is free once these MMAs have read it
          }
          __syncwarp();
          it_mma++;
        }
        if (elect_one()) tc_commit(done);  // accumulator of this pass complete
        __syncwarp();
      }
      if (tr) trace[ti++] = clock64();  // MMAs of the pass issued
      __syncwarp();
      mbar_wait(done, done_phase);
      done_phase ^= 1;
      __syncwarp();
      tc_fence_after();
      if (tr) trace[ti++] = clock64();  // accumulator complete

      const int mode = a.pass_mode[p];
      if (mode == PASS_SIM) {
        // similarity pass: 16 columns = h . (P W)_j; + P_j . b
        float v[16];
        tmem_ld16(t_lane, v);
#pragma unroll
        for (int j = 0; j < 16; j++) sim[j] = v[j] + __ldg(a.fold_b + j);
        tc_fence_before();
        __syncthreads();
        tc_fence_after();
        if (tr) trace[ti++] = clock64();
        continue;
      }
      const float* __restrict__ B = a.bias[l] + n0;
      if (mode == PASS_HIDDEN) {
        for (int c = 0; c < Np; c += 16) {
          float v[16];
          tmem_ld16(t_lane + c, v);
#pragma unroll
          for (int q4 = 0; q4 < 4; q4++) {
            const float4 b4 = __ldg(reinterpret_cast<const float4*>(B + c) + q4);
            v[q4 * 4 + 0] = fmaxf(v[q4 * 4 + 0] + b4.x, 0.0f);
            v[q4 * 4 + 1] = fmaxf(v[q4 * 4 + 1] + b4.y, 0.0f);
            v[q4 * 4 + 2] = fmaxf(v[q4 * 4 + 2] + b4.z, 0.0f);
            v[q4 * 4 + 3] = fmaxf(v[q4 * 4 + 3] + b4.w, 0.0f);
          }
          if constexpr (BX2) {
            uint4 hi, lo;
            bf16_split8(v, hi, lo);
            A[((c >> 3) + 0) * TC_M + tid] = hi;
            A_lo[((c >> 3) + 0) * TC_M + tid] = lo;
            bf16_split8(v + 8, hi, lo);
            A[((c >> 3) + 1) * TC_M + tid] = hi;
            A_lo[((c >> 3) + 1) * TC_M + tid] = lo;
          } else if constexpr (F16) {
            A[((c >> 3) + 0) * TC_M + tid] =
                make_uint4(pack_h2(v[0], v[1]), pack_h2(v[2], v[3]), pack_h2(v[4], v[5]), pack_h2(v[6], v[7]));
            A[((c >> 3) + 1) * TC_M + tid] =
                make_uint4(pack_h2(v[8], v[9]), pack_h2(v[10], v[11]), pack_h2(v[12], v[13]), pack_h2(v[14], v[15]));
          } else {
#pragma unroll
            for (int q4 = 0; q4 < 4; q4++)
              A[((c >> 2) + q4) * TC_M + tid] =
                  make_uint4(tf32_rn(v[q4 * 4 + 0]), tf32_rn(v[q4 * 4 + 1]), tf32_rn(v[q4 * 4 + 2]), tf32_rn(v[q4 * 4 + 3]));
          }
        }
        fence_async_smem();
      } else if (mode == PASS_NORM) {
        // last layer, first time round: only the sum of squares (nothing can be written before the norm is known)
        for (int c = 0; c < Np; c += 16) {
          float v[16];
          tmem_ld16(t_lane + c, v);
#pragma unroll
          for (int q4 = 0; q4 < 4; q4++) {
            const float4 b4 = __ldg(reinterpret_cast<const float4*>(B + c) + q4);
            v[q4 * 4 + 0] += b4.x; v[q4 * 4 + 1] += b4.y; v[q4 * 4 + 2] += b4.z; v[q4 * 4 + 3] += b4.w;
          }
#pragma unroll
          for (int i = 0; i < 16; i++) ssp[i & 3] = fmaf(v[i], v[i], ssp[i & 3]);
        }
        if (p == a.last_norm_pass) {
          norm = sqrtf((ssp[0] + ssp[1]) + (ssp[2] + ssp[3]));
          if (n_ph && live) {
            float max_neg = -INFINITY;
            const float inv = 1.0f / norm;
#pragma unroll
            for (int j = 0; j < 16; j++)
              if (j >= n_pos && j < n_ph) max_neg = fmaxf(max_neg, sim[j] * inv);
#pragma unroll
            for (int j = 0; j < 16; j++)
              if (j < n_pos) relev[(size_t)j * n + pix] = pair_softmax_pos(sim[j] * inv, max_neg);
          }
        }
      } else {
        // PASS_STORE: the last layer's MMAs once more (cheaper than parking 2 KB per pixel in L2 and rescaling it:
        // measured 22.9 GB of L2 traffic for 3.2 GB of output that way), normalised rows straight to `out`
        for (int c = 0; c < Np; c += 16) {
          float v[16];
          tmem_ld16(t_lane + c, v);
          if (live) {
#pragma unroll
            for (int q4 = 0; q4 < 4; q4++) {
              const float4 b4 = __ldg(reinterpret_cast<const float4*>(B + c) + q4);
              *reinterpret_cast<float4*>(out + pix * NL + n0 + c + q4 * 4) =
                  make_float4((v[q4 * 4 + 0] + b4.x) / norm, (v[q4 * 4 + 1] + b4.y) / norm,
                              (v[q4 * 4 + 2] + b4.z) / norm, (v[q4 * 4 + 3] + b4.w) / norm);
            }
          }
        }
      }
      tc_fence_before();
      __syncthreads();
      tc_fence_after();
      if (tr) trace[ti++] = clock64();  // epilogue done
    }
  }
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TC_TMEM_COLS) : "memory");
}

long long* g_dec_trace = nullptr;  // debug: device buffer of clock64 stamps (lsr_debug_decode_trace)

int sm_count() {  // of the current device
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
      n <= 0)
    n = 148;
  return n;
}

// passes of a decoder and the element offset of each in the packed blob (in 16-byte units)
struct PassPlan {
  int n;
  int layer[TC_MAX_PASSES], n0[TC_MAX_PASSES], np[TC_MAX_PASSES], kp[TC_MAX_PASSES];
  size_t off16[TC_MAX_PASSES];
  size_t total16;
};
PassPlan plan_passes(const lsr_decoder* d, int mode) {
  PassPlan pl{};
  const int E = mode == TC_TF32 ? 4 : 8, parts = mode == TC_BX2 ? 2 : 1;
  size_t off = 0;
  for (int l = 0; l < d->n_layers; l++) {
    const int K = d->dims[l], N = d->dims[l + 1], Kp = (K + 2 * E - 1) & ~(2 * E - 1);
    const int halves = N > 256 ? 2 : 1;
    for (int h = 0; h < halves; h++) {
      const int i = pl.n++;
      pl.layer[i] = l; pl.np[i] = N / halves; pl.n0[i] = h * (N / halves); pl.kp[i] = Kp; pl.off16[i] = off;
      off += (size_t)(Kp / E) * (N / halves) * parts;
    }
  }
  pl.total16 = off;
  return pl;
}

inline int tc_mode_of(int precision) {
  return precision == LSR_DECODE_F16 ? TC_F16 : (precision == LSR_DECODE_BF16X2 ? TC_BX2 : TC_TF32);
}
inline const void* packed_of(const lsr_decoder* d, int precision) {
  return precision == LSR_DECODE_F16 ? d->packed_f16 : (precision == LSR_DECODE_BF16X2 ? d->packed_bf16x2 : d->packed_tf32);
}

template <int MODE>
cudaError_t launch_tc(const lsr_decoder* d, const float* in, int64_t n, float* out, const float* phrases, int n_pos,
                      int n_neg, float* relev, void* work, cudaStream_t s) {
  using C = TcCfg<MODE>;
  TcArgs a{};
  a.n_layers = d->n_layers;
  for (int l = 0; l <= d->n_layers; l++) a.dims[l] = d->dims[l];
  for (int l = 0; l < d->n_layers; l++) a.bias[l] = d->bias[l];
  a.packed = MODE == TC_F16 ? d->packed_f16 : (MODE == TC_BX2 ? d->packed_bf16x2 : d->packed_tf32);
  const PassPlan pl = plan_passes(d, MODE);
  const int last = d->n_layers - 1, K7 = d->dims[last], K7p = (K7 + 2 * C::E - 1) & ~(2 * C::E - 1);
  if (phrases) {  // PW and P b into the work buffer: [K7p/E][16] 16-byte chunks, then 16 floats
    uint4* pw = reinterpret_cast<uint4*>(work);
    float* pb = reinterpret_cast<float*>(pw + (size_t)(K7p / C::E) * 16 * C::PARTS);
    const int items = (K7p / C::E) * 16 + 16;
    fold_phrases_kernel<<<(items + 63) / 64, 64, 0, s>>>(phrases, n_pos + n_neg, d->weight[last], d->bias[last],
                                                        d->dims[d->n_layers], K7, K7p, MODE, pw, pb);
    count_launch();
    a.fold = pw;
    a.fold_b = pb;
  }
  int ns = 0, np = 0;
  auto add_pass = [&](int layer, int n0, int Np, int Kp, size_t off16, int mode) -> bool {
    if (np >= TC_MAX_PASSES) return false;
    a.pass_layer[np] = layer; a.pass_n0[np] = n0; a.pass_np[np] = Np; a.pass_mode[np] = mode;
    const int chunks = Kp / C::E;
    int per = (C::STAGE_BYTES / (Np * 16 * C::PARTS)) & ~1;
    if (per < 2) per = 2;
    for (int c = 0; c < chunks; c += per) {
      if (ns >= TC_MAX_SLICES) return false;
      TcSlice& q = a.sl[ns++];
      q.off16 = (uint32_t)(off16 + (size_t)c * Np * C::PARTS);
      q.nch = (uint16_t)(chunks - c < per ? chunks - c : per);
      q.c0 = (uint8_t)c;
      q.pass = (uint8_t)np;
    }
    a.pass_slice_end[np++] = ns;
    return true;
  };
  bool sim_done = !phrases;
  for (int p = 0; p < pl.n; p++) {
    if (!sim_done && pl.layer[p] == last) {  // the similarity pass reads the last layer's input: right before it
      if (!add_pass(d->n_layers, 0, 16, K7p, 0, PASS_SIM)) return cudaErrorInvalidValue;
      sim_done = true;
    }
    if (!add_pass(pl.layer[p], pl.n0[p], pl.np[p], pl.kp[p], pl.off16[p], pl.layer[p] == last ? PASS_NORM : PASS_HIDDEN))
      return cudaErrorInvalidValue;
  }
  a.last_norm_pass = np - 1;
  if (out)
    for (int p = 0; p < pl.n; p++)
      if (pl.layer[p] == last && !add_pass(last, pl.n0[p], pl.np[p], pl.kp[p], pl.off16[p], PASS_STORE))
        return cudaErrorInvalidValue;
  a.n_passes = np;
  a.n_slices = ns;
  // Two co-resident CTAs that start together stay in phase (both in their MMA phase, then both in their epilogue) and
  // share the tensor pipe and the TMEM read port; a one-off delay of the second wave of CTAs puts one CTA's epilogues
  // under the other's MMAs.  LSR_DEC_STAGGER (cycles) overrides for experiments.
  if (C::CTAS_PER_SM > 1) {
    static int stagger = -1;
    if (stagger < 0) {
      const char* e = getenv("LSR_DEC_STAGGER");
      stagger = e ? atoi(e) : TC_F16_STAGGER;
      if (stagger < 0) stagger = 0;
    }
    a.stagger = stagger;
  }
  {  // per launch: the attribute is per device, and a process may use several
    cudaError_t e = cudaFuncSetAttribute(decode_tc_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
    if (e != cudaSuccess) return e;
  }
  const int64_t tiles = (n + TC_M - 1) / TC_M, slots = (int64_t)sm_count() * C::CTAS_PER_SM;
  const int grid = (int)(tiles < slots ? tiles : slots);
  decode_tc_kernel<MODE><<<grid, TC_M, C::SMEM, s>>>(a, in, n, out, phrases, n_pos, n_neg, relev, g_dec_trace);
  count_launch();
  return cudaGetLastError();
}

}  // namespace

void decoder_set_trace(void* dev_buf) { g_dec_trace = reinterpret_cast<long long*>(dev_buf); }

size_t decoder_packed_bytes(const lsr_decoder* d, int precision) {
  return plan_passes(d, tc_mode_of(precision)).total16 * 16;
}

void launch_decoder_pack(const lsr_decoder* d, int precision, void* packed, cudaStream_t s) {
  const int mode = tc_mode_of(precision);
  const PassPlan pl = plan_passes(d, mode);
  for (int p = 0; p < pl.n; p++) {
    const int l = pl.layer[p], K = d->dims[l];
    const int items = (pl.kp[p] / (mode == TC_TF32 ? 4 : 8)) * pl.np[p];
    pack_weights_kernel<<<(items + 255) / 256, 256, 0, s>>>(d->weight[l] + (size_t)pl.n0[p] * K, K, pl.np[p], pl.kp[p],
                                                            mode, reinterpret_cast<uint4*>(packed) + pl.off16[p]);
    count_launch();
  }
}

// Why a tensor-core variant cannot take this decoder (null if it can).
const char* decoder_tc_unsupported(const lsr_decoder* d, int precision) {
  const void* packed = packed_of(d, precision);
  if (!packed) return "the packed weights of this precision are null (call lsr_decoder_pack first)";
  if ((uintptr_t)packed & 15) return "packed weights must be 16-byte aligned";
  for (int l = 0; l < d->n_layers; l++) {
    const int K = d->dims[l], N = d->dims[l + 1];
    if (K > TC_MAXK) return "a layer input wider than 256";
    if (N % 16 || (N > 256 && N % 32)) return "layer widths must be multiples of 16 (of 32 above 256)";
    if ((uintptr_t)d->bias[l] & 15) return "bias pointers must be 16-byte aligned";
  }
  return nullptr;
}

cudaError_t launch_decode(const lsr_decoder* d, const float* in, int64_t n, int precision, float* out,
                          const float* phrases, int n_pos, int n_neg, float* relev, void* work, cudaStream_t s) {
  if (precision == LSR_DECODE_TF32) return launch_tc<TC_TF32>(d, in, n, out, phrases, n_pos, n_neg, relev, work, s);
  if (precision == LSR_DECODE_F16) return launch_tc<TC_F16>(d, in, n, out, phrases, n_pos, n_neg, relev, work, s);
  if (precision == LSR_DECODE_BF16X2) return launch_tc<TC_BX2>(d, in, n, out, phrases, n_pos, n_neg, relev, work, s);
  DecArgs a{};
  a.n_layers = d->n_layers;
  for (int l = 0; l <= d->n_layers; l++) a.dims[l] = d->dims[l];
  for (int l = 0; l < d->n_layers; l++) { a.weight[l] = d->weight[l]; a.bias[l] = d->bias[l]; }
  const int smem = 2 * F_TP * F_LD * 4;
  {
    cudaError_t e = cudaFuncSetAttribute(decode_fp32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
  }
  const int64_t tiles = (n + F_TP - 1) / F_TP;
  const int grid = (int)(tiles < (int64_t)sm_count() * 4 ? tiles : (int64_t)sm_count() * 4);
  decode_fp32_kernel<<<grid, F_THREADS, smem, s>>>(a, in, n, out, phrases, n_pos, n_neg, relev);
  count_launch();
  return cudaGetLastError();
}

}  // namespace lsr
