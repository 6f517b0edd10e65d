// decoder.cu — eval-side decode of rendered language features: the autoencoder's decoder MLP
// (autoencoder/model.py:18-25,42-46: Linear, (ReLU, Linear)*, x / ||x||; 3 -> 16 -> 32 -> 64 -> 128 -> 256 -> 256 -> 512
// with the defaults of eval/evaluate_iou_loc.py) followed by the CLIP relevancy of every pixel against the phrase
// embeddings (eval/openclip_encoder.py:42-56, 96-111).  This is the one dense contraction next to the rasterizer.
// The reference runs seven cuBLAS SGEMMs with [N,16..512] fp32 activations round-tripping HBM (about 10 KB per pixel)
// and then an [N,512] x [512,n] mm per level; here ONE kernel keeps a tile of pixels on chip from the 3-d input to
// the relevancy value: 12 B read and 4*n_pos B written per pixel (plus 2 KB if the restored feature is wanted).
//
// Two kernels behind lsr_decode:
//   * decode_tc_kernel (precision = LSR_DECODE_TF32): 128 pixels per CTA = the 128 TMEM lanes.  Every layer is a
//     tcgen05.mma.kind::tf32 chain (M = 128, N = layer width, K = 8 per instruction) issued by one thread; the A operand
//     (activations, fp32 in the canonical no-swizzle K-major core-matrix layout: [K/4][128 rows][4 floats]) stays in
//     shared memory, the B operand (weights, pre-packed once by lsr_decoder_pack into the same [K/4][N][4] layout)
//     streams from L2 through a 4-stage ring of 16 KB slices filled by cp.async.bulk (TMA 1-D) on mbarriers, the
//     accumulator lives in TMEM (up to 512 columns).  Epilogue per layer: tcgen05.ld, + bias, ReLU, back into the A
//     buffer; last layer: norm, phrase dot products and the softmax pair on the CUDA cores straight out of TMEM.
//   * decode_fp32_kernel (precision = LSR_DECODE_FP32): the same fusion on the FP32 pipe (32 pixels per CTA,
//     activations ping-pong in shared memory), IEEE fp32 like the reference's SGEMM — the exact-arithmetic variant.
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
// weight packing for the tensor-core variant: layer l -> [Kp/4][N][4] floats (Kp = K rounded up to 8, zero padded):
// the canonical no-swizzle K-major UMMA layout (8-row x 16-byte core matrices, 128 B apart along N, N*16 B along K).
__global__ void pack_weights_kernel(const float* __restrict__ W, int K, int N, int Kp, float4* __restrict__ dst) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (Kp / 4) * N) return;
  const int kc = i / N, o = i - kc * N;
  float v[4];
#pragma unroll
  for (int j = 0; j < 4; j++) {
    const int k = kc * 4 + j;
    v[j] = k < K ? __ldg(W + (size_t)o * K + k) : 0.0f;
  }
  dst[i] = make_float4(v[0], v[1], v[2], v[3]);
}

// ------------------------------------------------------------------------------------------------------------------
// tcgen05 variant
constexpr int TC_M = 128;
constexpr int TC_STAGES = 4;
constexpr int TC_STAGE_BYTES = 16384;
constexpr int TC_MAXK = 256;
constexpr int TC_A_BYTES = TC_M * TC_MAXK * 4;                         // 128 KB
constexpr int TC_PH_BYTES = DEC_MAXPH * DEC_MAXW * 4;                  // 32 KB
constexpr int TC_BAR_OFF = TC_A_BYTES + TC_STAGES * TC_STAGE_BYTES + TC_PH_BYTES;
constexpr int TC_SMEM = TC_BAR_OFF + 128;
constexpr int TC_MAX_SLICES = 96;
constexpr int TC_TMEM_COLS = 512;

struct TcSlice {
  uint32_t off16;  // offset of the slice in the packed blob, 16-byte units
  uint16_t nch;    // 16-byte K chunks in the slice (even)
  uint8_t c0;      // first chunk index within the layer
  uint8_t layer;
};
struct TcArgs {
  int n_layers;
  int dims[DEC_MAXL + 1];
  const float* bias[DEC_MAXL];
  const float* packed;
  int n_slices;
  int slice_end[DEC_MAXL];  // one past the last slice of each layer
  int variant;              // debug: bit 0 swaps the LBO / SBO fields of the shared-memory descriptors
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
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
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

__global__ void __launch_bounds__(TC_M, 1)
decode_tc_kernel(const __grid_constant__ TcArgs a, const float* __restrict__ in, int64_t n, float* __restrict__ out,
                 const float* __restrict__ phrases, int n_pos, int n_neg, float* __restrict__ relev) {
  extern __shared__ __align__(1024) uint8_t tsm[];
  float4* A = reinterpret_cast<float4*>(tsm);
  uint8_t* stages = tsm + TC_A_BYTES;
  float* ph = reinterpret_cast<float*>(stages + TC_STAGES * TC_STAGE_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(tsm + TC_BAR_OFF);  // full[4], empty[4], done
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * TC_STAGES + 1);
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + TC_STAGES), done = smem_u32(bars + 2 * TC_STAGES);
  const uint32_t a_base = smem_u32(A), st_base = smem_u32(stages);
  const int NL = a.dims[a.n_layers];
  const int n_ph = phrases ? n_pos + n_neg : 0;

  if (tid == 0) {
    for (int i = 0; i < TC_STAGES; i++) { mbar_init(full0 + 8 * i, 1); mbar_init(empty0 + 8 * i, 1); }
    mbar_init(done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "n"(TC_TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int i = tid; i < n_ph * NL; i += TC_M) ph[i] = __ldg(phrases + i);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t t_lane = tmem + ((uint32_t)(warp * 32) << 16);  // this warp's quarter of the 128 lanes

  uint32_t it_load = 0, it_mma = 0;  // thread 0: slices loaded / multiplied so far (ring position and phase)
  uint32_t done_phase = 0;
  const int64_t ntiles = (n + TC_M - 1) / TC_M;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t pix = tile * TC_M + tid;
    const bool live = pix < n;
    {  // input row -> A, K padded to 8 with zeros
      const int d0 = a.dims[0], kc_n = ((d0 + 7) & ~7) >> 2;
      for (int kc = 0; kc < kc_n; kc++) {
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const int k = kc * 4 + j;
          v[j] = (live && k < d0) ? __ldg(in + pix * d0 + k) : 0.0f;
        }
        A[kc * TC_M + tid] = make_float4(v[0], v[1], v[2], v[3]);
      }
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    int next_load = 0;  // thread 0: next slice of this tile to request
    for (int l = 0; l < a.n_layers; l++) {
      const int N = a.dims[l + 1];
      if (tid == 0) {
        const int s_begin = l ? a.slice_end[l - 1] : 0, s_end = a.slice_end[l];
        const int n_mma = N > 256 ? 2 : 1, Nm = N / n_mma;
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(Nm >> 3) << 17) | ((TC_M >> 4) << 24);
        const uint32_t a_lbo = TC_M * 16, b_lbo = (uint32_t)N * 16, sbo = 128;
        for (int s = s_begin; s < s_end; s++) {
          while (next_load < a.n_slices && next_load <= s + 2) {  // keep up to three slices in flight
            const TcSlice& q = a.sl[next_load];
            const uint32_t st = it_load % TC_STAGES, use = it_load / TC_STAGES;
            mbar_wait(empty0 + 8 * st, (use & 1) ^ 1);
            const uint32_t bytes = (uint32_t)q.nch * (uint32_t)a.dims[q.layer + 1] * 16u;
            mbar_expect_tx(full0 + 8 * st, bytes);
            bulk_g2s(st_base + st * TC_STAGE_BYTES, reinterpret_cast<const uint8_t*>(a.packed) + (size_t)q.off16 * 16,
                     bytes, full0 + 8 * st);
            it_load++;
            next_load++;
          }
          const TcSlice& q = a.sl[s];
          const uint32_t st = it_mma % TC_STAGES, use = it_mma / TC_STAGES;
          mbar_wait(full0 + 8 * st, use & 1);
          tc_fence_after();
          const uint32_t b_base = st_base + st * TC_STAGE_BYTES;
          for (int j = 0; j < q.nch / 2; j++) {
            const uint32_t a_addr = a_base + (uint32_t)(q.c0 + 2 * j) * a_lbo;
            const uint64_t ad = (a.variant & 1) ? smem_desc(a_addr, sbo, a_lbo) : smem_desc(a_addr, a_lbo, sbo);
            for (int h = 0; h < n_mma; h++) {
              const uint32_t b_addr = b_base + (uint32_t)(2 * j) * b_lbo + (uint32_t)(h * Nm) * 16;
              const uint64_t bd = (a.variant & 1) ? smem_desc(b_addr, sbo, b_lbo) : smem_desc(b_addr, b_lbo, sbo);
              tc_mma_tf32(tmem + (uint32_t)(h * Nm), ad, bd, idesc, (q.c0 + j) > 0 ? 1u : 0u);
            }
          }
          tc_commit(empty0 + 8 * st);  // the slice's stage is free once these MMAs have read it
          it_mma++;
        }
        tc_commit(done);  // accumulator of this layer complete
      }
      __syncwarp();
      mbar_wait(done, done_phase);
      done_phase ^= 1;
      __syncwarp();
      tc_fence_after();

      const float* __restrict__ B = a.bias[l];
      if (l + 1 < a.n_layers) {
        for (int c = 0; c < N; c += 16) {
          float v[16];
          tmem_ld16(t_lane + c, v);
#pragma unroll
          for (int q4 = 0; q4 < 4; q4++) {
            float4 o;
            o.x = fmaxf(v[q4 * 4 + 0] + __ldg(B + c + q4 * 4 + 0), 0.0f);
            o.y = fmaxf(v[q4 * 4 + 1] + __ldg(B + c + q4 * 4 + 1), 0.0f);
            o.z = fmaxf(v[q4 * 4 + 2] + __ldg(B + c + q4 * 4 + 2), 0.0f);
            o.w = fmaxf(v[q4 * 4 + 3] + __ldg(B + c + q4 * 4 + 3), 0.0f);
            A[((c >> 2) + q4) * TC_M + tid] = o;
          }
        }
        fence_async_smem();
      } else {
        float ss = 0.0f, sim[DEC_MAXPH];
#pragma unroll
        for (int j = 0; j < DEC_MAXPH; j++) sim[j] = 0.0f;
        for (int c = 0; c < N; c += 16) {
          float v[16];
          tmem_ld16(t_lane + c, v);
#pragma unroll
          for (int i = 0; i < 16; i++) {
            v[i] += __ldg(B + c + i);
            ss = fmaf(v[i], v[i], ss);
          }
#pragma unroll
          for (int j = 0; j < DEC_MAXPH; j++) {
            if (j < n_ph) {
              const float4* pj = reinterpret_cast<const float4*>(ph + j * NL + c);
#pragma unroll
              for (int q4 = 0; q4 < 4; q4++) {
                const float4 w = pj[q4];
                sim[j] = fmaf(v[q4 * 4 + 0], w.x, sim[j]);
                sim[j] = fmaf(v[q4 * 4 + 1], w.y, sim[j]);
                sim[j] = fmaf(v[q4 * 4 + 2], w.z, sim[j]);
                sim[j] = fmaf(v[q4 * 4 + 3], w.w, sim[j]);
              }
            }
          }
        }
        const float norm = sqrtf(ss);
        if (n_ph && live) {
          float max_neg = -INFINITY;
#pragma unroll
          for (int j = 0; j < DEC_MAXPH; j++)
            if (j >= n_pos && j < n_ph) max_neg = fmaxf(max_neg, sim[j] / norm);
#pragma unroll
          for (int j = 0; j < DEC_MAXPH; j++)
            if (j < n_pos) relev[(size_t)j * n + pix] = pair_softmax_pos(sim[j] / norm, max_neg);
        }
        if (out) {
          for (int c = 0; c < N; c += 16) {
            float v[16];
            tmem_ld16(t_lane + c, v);  // whole warp takes part (.sync.aligned); only live rows store
            if (live) {
#pragma unroll
              for (int q4 = 0; q4 < 4; q4++) {
                float4 o;
                o.x = (v[q4 * 4 + 0] + __ldg(B + c + q4 * 4 + 0)) / norm;
                o.y = (v[q4 * 4 + 1] + __ldg(B + c + q4 * 4 + 1)) / norm;
                o.z = (v[q4 * 4 + 2] + __ldg(B + c + q4 * 4 + 2)) / norm;
                o.w = (v[q4 * 4 + 3] + __ldg(B + c + q4 * 4 + 3)) / norm;
                *reinterpret_cast<float4*>(out + pix * N + c + q4 * 4) = o;
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncthreads();
      tc_fence_after();
    }
  }
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TC_TMEM_COLS) : "memory");
}

int g_sm_count = 0;
int sm_count() {
  if (!g_sm_count) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (g_sm_count <= 0) g_sm_count = 148;
  }
  return g_sm_count;
}

size_t packed_floats(const lsr_decoder* d, size_t* layer_off) {
  size_t off = 0;
  for (int l = 0; l < d->n_layers; l++) {
    if (layer_off) layer_off[l] = off;
    off += (size_t)((d->dims[l] + 7) & ~7) * d->dims[l + 1];
  }
  return off;
}

}  // namespace

size_t decoder_packed_bytes(const lsr_decoder* d) { return packed_floats(d, nullptr) * 4; }

void launch_decoder_pack(const lsr_decoder* d, float* packed, cudaStream_t s) {
  size_t off[DEC_MAXL];
  packed_floats(d, off);
  for (int l = 0; l < d->n_layers; l++) {
    const int K = d->dims[l], N = d->dims[l + 1], Kp = (K + 7) & ~7;
    const int items = (Kp / 4) * N;
    pack_weights_kernel<<<(items + 255) / 256, 256, 0, s>>>(d->weight[l], K, N, Kp,
                                                            reinterpret_cast<float4*>(packed + off[l]));
    count_launch();
  }
}

// Why the tensor-core variant cannot take this decoder (null if it can).
const char* decoder_tc_unsupported(const lsr_decoder* d) {
  if (!d->packed) return "lsr_decoder.packed is null (call lsr_decoder_pack first)";
  if ((uintptr_t)d->packed & 15) return "lsr_decoder.packed must be 16-byte aligned";
  for (int l = 0; l < d->n_layers; l++) {
    const int K = d->dims[l], N = d->dims[l + 1];
    if (K > TC_MAXK) return "a layer input wider than 256";
    if (N % 16 || (N > 256 && N % 32)) return "layer widths must be multiples of 16 (of 32 above 256)";
    if (2 * N * 16 > TC_STAGE_BYTES) return "layer wider than 512";
  }
  return nullptr;
}

cudaError_t launch_decode(const lsr_decoder* d, const float* in, int64_t n, int precision, float* out,
                          const float* phrases, int n_pos, int n_neg, float* relev, cudaStream_t s) {
  if (precision == LSR_DECODE_FP32) {
    DecArgs a{};
    a.n_layers = d->n_layers;
    for (int l = 0; l <= d->n_layers; l++) a.dims[l] = d->dims[l];
    for (int l = 0; l < d->n_layers; l++) { a.weight[l] = d->weight[l]; a.bias[l] = d->bias[l]; }
    const int smem = 2 * F_TP * F_LD * 4;
    static bool attr = false;
    if (!attr) {
      cudaError_t e = cudaFuncSetAttribute(decode_fp32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      if (e != cudaSuccess) return e;
      attr = true;
    }
    const int64_t tiles = (n + F_TP - 1) / F_TP;
    const int grid = (int)(tiles < (int64_t)sm_count() * 4 ? tiles : (int64_t)sm_count() * 4);
    decode_fp32_kernel<<<grid, F_THREADS, smem, s>>>(a, in, n, out, phrases, n_pos, n_neg, relev);
    count_launch();
    return cudaGetLastError();
  }
  TcArgs a{};
  a.n_layers = d->n_layers;
  for (int l = 0; l <= d->n_layers; l++) a.dims[l] = d->dims[l];
  for (int l = 0; l < d->n_layers; l++) a.bias[l] = d->bias[l];
  a.packed = d->packed;
  const char* v = getenv("LSR_TC_VARIANT");
  a.variant = v ? atoi(v) : 0;
  size_t off[DEC_MAXL];
  packed_floats(d, off);
  int ns = 0;
  for (int l = 0; l < d->n_layers; l++) {
    const int Kp = (d->dims[l] + 7) & ~7, N = d->dims[l + 1], chunks = Kp / 4;
    int per = (TC_STAGE_BYTES / (N * 16)) & ~1;
    if (per < 2) per = 2;
    for (int c = 0; c < chunks; c += per) {
      if (ns >= TC_MAX_SLICES) return cudaErrorInvalidValue;
      TcSlice& q = a.sl[ns++];
      q.off16 = (uint32_t)((off[l] * 4 + (size_t)c * N * 16) / 16);
      q.nch = (uint16_t)(chunks - c < per ? chunks - c : per);
      q.c0 = (uint8_t)c;
      q.layer = (uint8_t)l;
    }
    a.slice_end[l] = ns;
  }
  a.n_slices = ns;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(decode_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM);
    if (e != cudaSuccess) return e;
    attr = true;
  }
  const int64_t tiles = (n + TC_M - 1) / TC_M;
  const int grid = (int)(tiles < (int64_t)sm_count() ? tiles : (int64_t)sm_count());
  decode_tc_kernel<<<grid, TC_M, TC_SMEM, s>>>(a, in, n, out, phrases, n_pos, n_neg, relev);
  count_launch();
  return cudaGetLastError();
}

}  // namespace lsr
