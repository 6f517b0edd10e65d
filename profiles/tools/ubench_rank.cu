// Micro-benchmark (analysis helper): cost of three ways to find the lanes holding the same 8-bit digit, per warp-item,
// with 32 resident warps per SM: (a) 8 ballots, (b) match.any, (c) shared-memory atomicOr + read-back.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t match_ballot(uint32_t d) {
  uint32_t peers = 0xffffffffu;
#pragma unroll
  for (int b = 0; b < 8; b++) {
    uint32_t m;
    asm("{\n\t.reg .pred p;\n\t.reg .b32 t;\n\tand.b32 t, %1, %2;\n\tsetp.ne.u32 p, t, 0;\n\t"
        "vote.sync.ballot.b32 %0, p, 0xffffffff;\n\t@!p not.b32 %0, %0;\n\t}" : "=r"(m) : "r"(d), "r"(1u << b));
    peers &= m;
  }
  return peers;
}
template <int MODE>
__global__ void __launch_bounds__(256) k(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int iters, int spread) {
  __shared__ uint32_t wmask[8][256];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int j = tid; j < 8 * 256; j += 256) (&wmask[0][0])[j] = 0;
  __syncthreads();
  uint32_t x = in[blockIdx.x * 256 + tid], acc = 0;
  for (int i = 0; i < iters; i++) {
    x = x * 1664525u + 1013904223u;
    uint32_t d;
    if (spread == 1) d = (x >> 13) & 0xffu;             // random 8-bit
    else if (spread == 0) d = (x >> 13) & 0x3u;         // 4 values
    else if (spread == 2) d = (lane + i) & 0xffu;       // consecutive integers
    else if (spread == 3) d = ((lane / 3) * 120 + lane % 3 + i * 7) & 0xffu;  // tile-id like
    else d = (i & 0xff);                                // all equal
    uint32_t peers;
    if (MODE == 0) peers = match_ballot(d);
    else if (MODE == 1) peers = __match_any_sync(0xffffffffu, d);
    else if (MODE == 3) {
      peers = __match_any_sync(0xffffffffu, d);
      if ((peers & ((1u << lane) - 1)) == 0) wmask[warp][d] += __popc(peers);
      __syncwarp();
    } else if (MODE == 4) {
      peers = __match_any_sync(0xffffffffu, d);
      peers += __reduce_add_sync(peers, x & 7u);
    } else if (MODE == 5) {
      peers = __match_any_sync(0xffffffffu, d);
      const uint32_t prev = wmask[warp][d];
      __syncwarp();
      if ((peers & ((1u << lane) - 1)) == 0) wmask[warp][d] = prev + __popc(peers);
      peers += prev;
      __syncwarp();
    } else {
      atomicOr(&wmask[warp][d], 1u << lane);
      __syncwarp();
      peers = wmask[warp][d];
      __syncwarp();
      if ((peers & ((1u << lane) - 1)) == 0) wmask[warp][d] = 0;
    }
    acc += __popc(peers);
  }
  out[blockIdx.x * 256 + tid] = acc;
}
int main() {
  const int blocks = 148 * 4, iters = 4096;
  uint32_t *in, *out;
  cudaMalloc(&in, blocks * 256 * 4); cudaMalloc(&out, blocks * 256 * 4);
  cudaMemset(in, 1, blocks * 256 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int spread = 4; spread >= 0; spread--)
    for (int mode = 1; mode < 6; mode += 4) {
      float best = 1e9;
      for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        if (mode == 0) k<0><<<blocks, 256>>>(in, out, iters, spread);
        if (mode == 1) k<1><<<blocks, 256>>>(in, out, iters, spread);
        if (mode == 2) k<2><<<blocks, 256>>>(in, out, iters, spread);
        if (mode == 3) k<3><<<blocks, 256>>>(in, out, iters, spread);
        if (mode == 4) k<4><<<blocks, 256>>>(in, out, iters, spread);
        if (mode == 5) k<5><<<blocks, 256>>>(in, out, iters, spread);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best;
      }
      const double warp_items = (double)blocks * 8 * iters;
      printf("spread=%d mode=%d (%s): %.3f ms, %.1f SM-cycles per warp-item (at 1.965 GHz, 148 SMs)\n", spread, mode,
             mode == 0 ? "8 ballots" : mode == 1 ? "match.any" : mode == 2 ? "smem atomicOr" : mode == 3 ? "match+leader RMW" : mode == 4 ? "match+redux(peers)" : "match+read-all+leader write", best,
             best * 1e-3 * 1.965e9 * 148 / warp_items);
    }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
