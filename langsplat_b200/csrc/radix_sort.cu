// radix_sort.cu — K4': onesweep-style stable LSD radix sort of (uint64 key, uint32 value) pairs.
// Replaces cub::DeviceRadixSort::SortPairs (RAST/cuda_rasterizer/rasterizer_impl.cu:306-311).  A stable ascending
// sort on bits [0, end_bit) is fully specified, so any correct implementation yields the reference's exact
// permutation.
//
// Structure: one histogram kernel computes the digit histograms of every pass in a single read of the keys; a tiny
// kernel turns them into exclusive bin bases; then one kernel per 8-bit digit does rank + scatter in a single sweep,
// chaining CTA prefixes through global memory with decoupled look-back (no separate scan pass between digits).
// Ranking inside a CTA is a warp-level multi-split: __match_any_sync groups equal digits, the group leader bumps a
// per-warp shared histogram, ranks come from popc of the lower-lane mask.  Keys and values are then staged in shared
// memory in CTA-sorted order so the global scatter is written as long coalesced runs.
// HBM-bound: algorithmic bytes per element 8 (histogram read) + 24 per pass.
#include <algorithm>
#include "common.cuh"

namespace lsr {

namespace {

constexpr uint32_t LB_AGG = 1u << 30, LB_INC = 2u << 30, LB_FLAG = 3u << 30, LB_VAL = ~LB_FLAG;
constexpr int WARPS = SORT_THREADS / 32;
constexpr size_t SORT_SMEM = (size_t)SORT_TILE * 12 + (size_t)WARPS * SORT_BINS * 4 + 2 * SORT_BINS * 4;

__global__ void __launch_bounds__(256)
sort_hist_kernel(const uint64_t* __restrict__ keys, uint32_t n, int passes, int end_bit, uint32_t* __restrict__ ghist) {
  __shared__ uint32_t sh[SORT_MAX_PASSES * SORT_BINS];
  for (int j = threadIdx.x; j < passes * SORT_BINS; j += 256) sh[j] = 0;
  __syncthreads();
  const uint32_t stride = gridDim.x * 256;
  const uint32_t n_round = (n + 31) & ~31u;  // keep warps converged for match_any
  for (uint32_t i = blockIdx.x * 256 + threadIdx.x; i < n_round; i += stride) {
    const bool ok = i < n;
    const uint64_t key = ok ? __ldg(keys + i) : 0ull;
    const uint32_t act = __ballot_sync(0xffffffffu, ok);
    if (!ok) continue;
    for (int p = 0; p < passes; p++) {
      const int shift = p * SORT_RADIX_BITS;
      const int bits = min(SORT_RADIX_BITS, end_bit - shift);
      const uint32_t d = (uint32_t)(key >> shift) & ((1u << bits) - 1);
      const uint32_t peers = __match_any_sync(act, d);
      if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&sh[p * SORT_BINS + d], __popc(peers));
    }
  }
  __syncthreads();
  for (int j = threadIdx.x; j < passes * SORT_BINS; j += 256)
    if (sh[j]) atomicAdd(&ghist[j], sh[j]);
}

// exclusive scan of each pass's 256-bin histogram, in place (one CTA per pass)
__global__ void __launch_bounds__(SORT_BINS) sort_scan_kernel(uint32_t* __restrict__ ghist) {
  __shared__ uint32_t s_w[SORT_BINS / 32];
  uint32_t* h = ghist + blockIdx.x * SORT_BINS;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const uint32_t c = h[t];
  uint32_t w = c;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t x = __shfl_up_sync(0xffffffffu, w, d);
    if (lane >= d) w += x;
  }
  if (lane == 31) s_w[warp] = w;
  __syncthreads();
  uint32_t pre = 0;
#pragma unroll
  for (int i = 0; i < SORT_BINS / 32; i++)
    if (i < warp) pre += s_w[i];
  h[t] = pre + w - c;
}

__global__ void __launch_bounds__(SORT_THREADS)
onesweep_kernel(const uint64_t* __restrict__ kin, uint64_t* __restrict__ kout, const uint32_t* __restrict__ vin,
                uint32_t* __restrict__ vout, uint32_t n, uint32_t* __restrict__ ticket,
                volatile uint32_t* __restrict__ lookback, const uint32_t* __restrict__ bin_base, int shift,
                uint32_t mask) {
  extern __shared__ __align__(16) unsigned char smem[];
  uint64_t* s_keys = reinterpret_cast<uint64_t*>(smem);
  uint32_t* s_vals = reinterpret_cast<uint32_t*>(smem + (size_t)SORT_TILE * 8);
  uint32_t* s_whist = reinterpret_cast<uint32_t*>(smem + (size_t)SORT_TILE * 12);
  uint32_t* s_binstart = s_whist + WARPS * SORT_BINS;
  uint32_t* s_gofs = s_binstart + SORT_BINS;
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_wsum[WARPS];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(ticket, 1u);
#pragma unroll
  for (int j = tid; j < WARPS * SORT_BINS; j += SORT_THREADS) s_whist[j] = 0;
  __syncthreads();
  const uint32_t tile = s_tile;
  const uint32_t tile_base = tile * SORT_TILE;
  if (tile_base >= n) return;
  const uint32_t valid = min((uint32_t)SORT_TILE, n - tile_base);

  // ---- load (warp-striped: warp w owns a contiguous 32*ITEMS chunk; item i of lane l is element i*32+l) ----
  const uint32_t wbase = tile_base + warp * (32 * SORT_ITEMS);
  uint64_t key[SORT_ITEMS];
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; i++) {
    const uint32_t gi = wbase + i * 32 + lane;
    key[i] = gi < n ? __ldg(kin + gi) : ~0ull;
  }
  // ---- warp-level multi-split ranking (stable: items in ascending i, lanes in ascending order) ----
  uint32_t rank[SORT_ITEMS];
  uint32_t* wh = s_whist + warp * SORT_BINS;
  const uint32_t lt = (1u << lane) - 1;
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; i++) {
    const uint32_t d = (uint32_t)(key[i] >> shift) & mask;
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    const int leader = __ffs(peers) - 1;
    uint32_t prev = 0;
    if (lane == leader) {
      prev = wh[d];
      wh[d] = prev + __popc(peers);
    }
    prev = __shfl_sync(0xffffffffu, prev, leader);
    rank[i] = prev + __popc(peers & lt);
    __syncwarp();
  }
  __syncthreads();

  // ---- per bin (thread t == bin t): exclusive scan over warps, CTA bin scan, decoupled look-back ----
  {
    const int t = tid;
    uint32_t cnt = 0;
#pragma unroll
    for (int w = 0; w < WARPS; w++) {
      const uint32_t c = s_whist[w * SORT_BINS + t];
      s_whist[w * SORT_BINS + t] = cnt;
      cnt += c;
    }
    uint32_t inc = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      uint32_t x = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += x;
    }
    if (lane == 31) s_wsum[warp] = inc;
    __syncthreads();
    uint32_t pre = 0;
#pragma unroll
    for (int w = 0; w < WARPS; w++)
      if (w < warp) pre += s_wsum[w];
    const uint32_t bin_start = pre + inc - cnt;
    s_binstart[t] = bin_start;
    // padding keys (~0) of a partial last tile all carry digit == mask and sort to the very end of the CTA
    uint32_t vcnt = cnt;
    if ((uint32_t)t == mask) vcnt -= (SORT_TILE - valid);
    uint32_t prefix = 0;
    volatile uint32_t* mine = lookback + (size_t)tile * SORT_BINS + t;
    if (tile == 0) {
      *mine = LB_INC | vcnt;
    } else {
      *mine = LB_AGG | vcnt;
      int k = (int)tile - 1;
      while (true) {
        uint32_t v;
        do { v = lookback[(size_t)k * SORT_BINS + t]; } while ((v & LB_FLAG) == 0);
        prefix += v & LB_VAL;
        if ((v & LB_FLAG) == LB_INC) break;
        k--;
      }
      *mine = LB_INC | (prefix + vcnt);
    }
    s_gofs[t] = __ldg(bin_base + t) + prefix - bin_start;
  }
  __syncthreads();

  // ---- stage keys and values in CTA-sorted order ----
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; i++) {
    const uint32_t d = (uint32_t)(key[i] >> shift) & mask;
    const uint32_t pos = s_binstart[d] + wh[d] + rank[i];
    rank[i] = pos;
    s_keys[pos] = key[i];
  }
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; i++) {
    const uint32_t gi = wbase + i * 32 + lane;
    s_vals[rank[i]] = gi < n ? __ldg(vin + gi) : 0u;
  }
  __syncthreads();
  // ---- coalesced scatter ----
  for (uint32_t p = tid; p < valid; p += SORT_THREADS) {
    const uint64_t k = s_keys[p];
    const uint32_t d = (uint32_t)(k >> shift) & mask;
    const uint32_t g = s_gofs[d] + p;
    kout[g] = k;
    vout[g] = s_vals[p];
  }
}

}  // namespace

int sort_pairs(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
               uint64_t* keys_tmp, uint32_t* vals_tmp, int64_t n, int end_bit, void* work, cudaStream_t s) {
  if (n <= 0) return 0;
  if (end_bit < 1) end_bit = 1;
  if (end_bit > 64) end_bit = 64;
  const int passes = (end_bit + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS;
  SortWork w = sort_work_view(work, n);
  cudaMemsetAsync(work, 0, w.zero_bytes, s);
  const uint32_t nblocks = (uint32_t)((n + SORT_TILE - 1) / SORT_TILE);
  static bool attr_done = false;
  if (!attr_done) {
    cudaFuncSetAttribute(onesweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SORT_SMEM);
    attr_done = true;
  }
  const uint32_t hist_blocks = (uint32_t)std::min<int64_t>((n + 2047) / 2048, (int64_t)148 * 8);
  sort_hist_kernel<<<hist_blocks, 256, 0, s>>>(keys_in, (uint32_t)n, passes, end_bit, w.hist);
  sort_scan_kernel<<<passes, SORT_BINS, 0, s>>>(w.hist);
  count_launch(2);
  const uint64_t* ksrc = keys_in;
  const uint32_t* vsrc = vals_in;
  for (int p = 0; p < passes; p++) {
    const bool to_out = ((passes - 1 - p) & 1) == 0;
    uint64_t* kdst = to_out ? keys_out : keys_tmp;
    uint32_t* vdst = to_out ? vals_out : vals_tmp;
    const int shift = p * SORT_RADIX_BITS;
    const int bits = min(SORT_RADIX_BITS, end_bit - shift);
    onesweep_kernel<<<nblocks, SORT_THREADS, SORT_SMEM, s>>>(
        ksrc, kdst, vsrc, vdst, (uint32_t)n, w.tickets + p, w.lookback + (size_t)p * sort_nblocks(n) * SORT_BINS,
        w.hist + p * SORT_BINS, shift, (1u << bits) - 1);
    count_launch();
    ksrc = kdst;
    vsrc = vdst;
  }
  return passes;
}

}  // namespace lsr
