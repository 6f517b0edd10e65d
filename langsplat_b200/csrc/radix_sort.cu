// radix_sort.cu — K4': onesweep-style stable LSD radix sort of (uint64 key, uint32 value) pairs.
// Replaces cub::DeviceRadixSort::SortPairs (RAST/cuda_rasterizer/rasterizer_impl.cu:306-311).  A stable ascending
// sort on bits [0, end_bit) is fully specified, so any correct implementation yields the reference's exact
// permutation.
//
// Structure
//  * The digit histograms of every pass come from ONE upstream read: in the forward they are accumulated by the
//    duplication kernel while it emits the keys (binning.cu), so the sort never re-reads the keys for counting; the
//    stand-alone entry point (lsr_sort_pairs) runs sort_hist_kernel instead.
//  * One kernel per digit does rank + scatter in a single sweep, chaining tile prefixes through global memory with
//    decoupled look-back (no scan pass between digits).  Each CTA turns the raw global histogram into exclusive bin
//    bases itself (256-bin block scan), so there is no separate scan launch either.
//  * Ranking inside a CTA is a warp-level multi-split: 8 ballots group the lanes holding the same digit (measured on
//    B200, profiles/tools/ubench_rank.cu, SM-cycles per warp instruction: 8 ballots 14; MATCH.ANY 5 when all lanes
//    agree but 66 when all differ; a shared-memory atomic 34), the group's lowest lane bumps a per-warp shared
//    histogram with a plain read-modify-write, ranks come from popc of the lower-lane mask.  Keys, then values, are staged in
//    shared memory in CTA-sorted order so the global scatter is written as coalesced runs.
//  * Digits are taken from a COMPACTED view of the key, (hi << lo_bits) | (lo - lo_min): the forward knows the range
//    of the depth word (min / max of the float bits, reduced in the preprocess kernel), and depth bits minus their
//    minimum need ~25 bits instead of 32, so a 45-bit key sorts in 5 eight-bit passes instead of 6.  The transform is
//    order preserving and the keys in memory stay the reference's exact values.
// HBM-bound: algorithmic bytes per element 24 per pass (12 read + 12 written).
#include <algorithm>
#include <mutex>
#include "common.cuh"

// ---- scratch of the allocation-free entry point (lsr_sort_pairs_static, bottom of this file) ----
// Work area (histograms, tickets, look-back words) + the two ping-pong buffers for at most LSR_SORT_STATIC_MAX_N pairs:
// 14.8 MB that every GPU which loads the library holds until the process exits.  In HOST code this name is only a
// stand-in for the symbol; the device address comes from cudaGetSymbolAddress (static_sort_scratch() below).
constexpr int64_t kStaticSortMaxN = LSR_SORT_STATIC_MAX_N;
constexpr size_t kStaticAlign = 256;
constexpr size_t static_up(size_t v) { return (v + kStaticAlign - 1) / kStaticAlign * kStaticAlign; }
constexpr size_t kStaticNBlocks = (size_t)((kStaticSortMaxN + lsr::SORT_TILE - 1) / lsr::SORT_TILE) + 1;
constexpr size_t kStaticWorkBytes = static_up(lsr::SORT_MAX_PASSES * lsr::SORT_BINS * 4) + static_up(lsr::SORT_MAX_PASSES * 4) +
                                    static_up((size_t)lsr::SORT_MAX_PASSES * kStaticNBlocks * lsr::SORT_BINS * 4) + 256;
constexpr size_t kStaticKeysOfs = static_up(kStaticWorkBytes);
constexpr size_t kStaticValsOfs = kStaticKeysOfs + static_up((size_t)kStaticSortMaxN * 8);
constexpr size_t kStaticScratchBytes = kStaticValsOfs + static_up((size_t)kStaticSortMaxN * 4);
static __device__ __align__(256) unsigned char g_static_sort_scratch[kStaticScratchBytes];

namespace lsr {

namespace {

#ifndef LSR_SORT_LB_WIN
#define LSR_SORT_LB_WIN 4
#endif
constexpr int LB_WIN = LSR_SORT_LB_WIN;
constexpr uint32_t LB_AGG = 1u << 30, LB_INC = 2u << 30, LB_FLAG = 3u << 30, LB_VAL = ~LB_FLAG;
constexpr int WARPS = SORT_THREADS / 32;
constexpr int BITS = SORT_RADIX_BITS;
constexpr int BINS = SORT_BINS;
static_assert(BINS == SORT_THREADS, "one bin per thread in the look-back");
// keys (8 B) staged first, the same region is reused for the values; per-warp histograms; bin starts; global offsets
constexpr size_t sort_smem(int items) {
  return (size_t)SORT_THREADS * items * 8 + (size_t)WARPS * BINS * 4 + BINS * 4 + (size_t)SORT_THREADS * items;
}

struct DigitMap {  // compacted key = (hi << lo_bits) | (lo - lo_min)
  uint32_t lo_min;
  int lo_bits;
};
__device__ __forceinline__ uint32_t digit_of(uint64_t key, const DigitMap m, int shift, uint32_t mask) {
  const uint32_t lo = (uint32_t)key - m.lo_min;
  const uint64_t c = ((key >> 32) << m.lo_bits) | (uint64_t)lo;
  return (uint32_t)(c >> shift) & mask;
}

// Stand-alone histogram of all passes (lsr_sort_pairs and the forward's depth-order sort of the Gaussians; the
// instance sort's histograms come from the duplication kernel).  packed_shift >= 0: the words are packed
// (compacted key << id_bits | id) and packed_shift = id_bits.
// Warp-private shared histograms, one plain read-modify-write per group of equal digits (no shared atomics).
__global__ void __launch_bounds__(256)
sort_hist_kernel(const uint64_t* __restrict__ keys, uint32_t n, int passes, int total_bits, DigitMap dm,
                 int packed_shift, uint32_t* __restrict__ ghist) {
  extern __shared__ uint32_t sh_hist[];  // [8 warps][passes][BINS]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int j = tid; j < 8 * passes * BINS; j += 256) sh_hist[j] = 0;
  __syncthreads();
  uint32_t* wh = sh_hist + warp * passes * BINS;
  const uint32_t stride = gridDim.x * 256;
  const uint32_t n_round = (n + 31) & ~31u;  // keep warps converged for match.any
  for (uint32_t i = blockIdx.x * 256 + tid; i < n_round; i += stride) {
    const bool ok = i < n;
    const uint64_t key = ok ? __ldg(keys + i) : 0ull;
    for (int p = 0; p < passes; p++) {
      const int shift = p * BITS;
      const int bits = min(BITS, total_bits - shift);
      const uint32_t mask = (1u << bits) - 1;
      const uint32_t d = !ok ? 0xffffffffu
                             : (packed_shift >= 0 ? ((uint32_t)(key >> (packed_shift + shift)) & mask)
                                                  : digit_of(key, dm, shift, mask));
      const uint32_t peers = __match_any_sync(0xffffffffu, d);
      if (ok && (peers & ((1u << lane) - 1)) == 0) wh[p * BINS + d] += __popc(peers);
      __syncwarp();
    }
  }
  __syncthreads();
  for (int j = tid; j < passes * BINS; j += 256) {
    uint32_t c = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) c += sh_hist[w * passes * BINS + j];
    if (c) atomicAdd(&ghist[j], c);
  }
}

// Per-pass digit extraction with all shift amounts precomputed on the host (PTX shifts clamp at 32, so the terms
// that do not apply to a pass vanish):  d = (((lo - lo_min) >> s1) | (hi << s2) | (hi >> s3)) & mask.
struct PassDigit {
  uint32_t lo_min, s1, s2, s3, mask;
};
__device__ __forceinline__ uint32_t shr32(uint32_t v, uint32_t s) {
  uint32_t r;
  asm("shr.b32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(s));
  return r;
}
__device__ __forceinline__ uint32_t shl32(uint32_t v, uint32_t s) {
  uint32_t r;
  asm("shl.b32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(s));
  return r;
}
__device__ __forceinline__ uint32_t pass_digit(uint64_t key, const PassDigit& pd) {
  const uint32_t lo = (uint32_t)key - pd.lo_min, hi = (uint32_t)(key >> 32);
  return (shr32(lo, pd.s1) | shl32(hi, pd.s2) | shr32(hi, pd.s3)) & pd.mask;
}

struct SortDynArgs {  // per-launch view of SortDyn (common.cuh)
  const uint32_t* n_dev;
  const uint32_t* bits_dev;
  uint32_t shift;
  uint2* ranges;
  int tile_shift;
};

// Lanes of the warp holding the same digit, from NB ballots.  MATCH.ANY would be one instruction, but on B200 its cost
// grows with the number of DISTINCT values in the warp (profiles/tools/ubench_rank.cu: 5 SM-cycles per warp
// instruction when all lanes agree, 66 when all differ — and radix digits mostly differ); 8 ballots cost ~14.
template <int NB>
__device__ __forceinline__ uint32_t match_digit(uint32_t d) {
  uint32_t peers = 0xffffffffu;
#pragma unroll
  for (int b = 0; b < NB; b++) {
    uint32_t m;
    asm("{\n\t.reg .pred p;\n\t.reg .b32 t;\n\t"
        "and.b32 t, %1, %2;\n\t"
        "setp.ne.u32 p, t, 0;\n\t"
        "vote.sync.ballot.b32 %0, p, 0xffffffff;\n\t"
        "@!p not.b32 %0, %0;\n\t}"
        : "=r"(m)
        : "r"(d), "r"(1u << b));
    peers &= m;
  }
  return peers;
}

// Shared-memory staging index of CTA-sorted position p.  Keys whose digits are consecutive (instances of one Gaussian
// cover consecutive tiles) land ~SORT_TILE/BINS = 16 positions apart, i.e. in the same bank for 8-byte elements; the
// XOR swizzle spreads them (ncu: the tile-id pass ran at 99 us against 53 us for the depth passes without it).
__device__ __forceinline__ uint32_t sw(uint32_t p) { return p ^ ((p >> 4) & 15u); }

// PAIRS = true: (key, value) pairs in separate arrays (lsr_sort_pairs, and forwards whose key + id exceed 64 bits).
// PAIRS = false: the value rides in the low bits of the 64-bit word (binning.cu packs compacted key | Gaussian id), so
// a pass moves 8 instead of 12 bytes per element and the whole value phase disappears; the last pass additionally
// writes the ids out as the u32 point list (`ids_out`).
#ifndef LSR_SORT_MINB
#define LSR_SORT_MINB 4  // CTAs per SM the register allocation aims at (4: 64 registers with ~100 B of spills)
#endif
#ifndef LSR_SORT_RANK2
// 1: ranking as two loops (all ballots first; then only the lowest lane of a group of equal digits does the
// shared-memory read-modify-write and the others take its old value by shuffle).  Fewer shared-memory accesses and a
// shorter chain on paper, slower on the GPU (call r2aa, C2: instance sort 132 -> 147 us, depth ordering 67 -> 75 us;
// with 80 registers 143 / 71): the divergent leader branch and the extra shuffle cost more than the all-lane LDS.
#define LSR_SORT_RANK2 0
#endif
#ifndef LSR_DEPTH_NB
#define LSR_DEPTH_NB 8  // ranking of the depth-ordering passes: eight ballots (0 = MATCH.ANY: 77 vs 65 us for the
                       // three passes at C2, GPU call r2d: the low digits of a depth word mostly differ within a warp)
#endif
template <int NB, bool PAIRS, int ITEMS>
__global__ void __launch_bounds__(SORT_THREADS, LSR_SORT_MINB)
onesweep_kernel(const uint64_t* __restrict__ kin, uint64_t* __restrict__ kout, const uint32_t* __restrict__ vin,
                uint32_t* __restrict__ vout, uint32_t n_bound, uint32_t* __restrict__ ticket,
                volatile uint32_t* __restrict__ lookback, const uint32_t* __restrict__ ghist, PassDigit pd,
                uint32_t* __restrict__ ids_out, uint32_t id_mask, SortDynArgs dyn) {
  // Sizes that only the device knows (the forward never reads them back): the element count, and for the depth
  // ordering the width of the key — a pass whose digit lies above it has nothing to do, and the last pass that has
  // emits the ids.
  if (dyn.bits_dev != nullptr) {
    const uint32_t width = __ldg(dyn.bits_dev);
    if (dyn.shift >= width) return;
    if (dyn.shift + BITS < width) ids_out = nullptr;
  }
  const uint32_t n = dyn.n_dev != nullptr ? min(__ldg(dyn.n_dev), n_bound) : n_bound;
  extern __shared__ __align__(16) unsigned char smem[];
  uint64_t* s_keys = reinterpret_cast<uint64_t*>(smem);
  uint32_t* s_vals = reinterpret_cast<uint32_t*>(smem);  // reuses the key region after the keys are written out
  uint32_t* s_whist = reinterpret_cast<uint32_t*>(smem + (size_t)(SORT_THREADS * ITEMS) * 8);  // [WARPS][BINS]
  uint32_t* s_gofs = s_whist + WARPS * BINS;
  uint8_t* s_dig = reinterpret_cast<uint8_t*>(s_gofs + BINS);  // digit of the element at CTA-sorted position p
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_wsum[WARPS], s_wsum2[WARPS];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(ticket, 1u);
#pragma unroll
  for (int j = tid; j < WARPS * BINS; j += SORT_THREADS) s_whist[j] = 0;
  __syncthreads();
  const uint32_t tile = s_tile;
  const uint32_t tile_base = tile * (SORT_THREADS * ITEMS);
  if (tile_base >= n) return;
  const uint32_t valid = min((uint32_t)(SORT_THREADS * ITEMS), n - tile_base);
  const bool full = valid == (SORT_THREADS * ITEMS);

  // ---- load (warp-striped: warp w owns a contiguous 32*ITEMS chunk; item i of lane l is element i*32+l) ----
  const uint32_t wofs = warp * (32 * ITEMS) + lane;  // + i*32 = position inside the tile
  const uint64_t* kp = kin + tile_base + wofs;
  uint64_t key[ITEMS];
  if (full) {
#pragma unroll
    for (int i = 0; i < ITEMS; i++) key[i] = __ldg(kp + i * 32);
  } else {
#pragma unroll
    for (int i = 0; i < ITEMS; i++) key[i] = (wofs + i * 32 < valid) ? __ldg(kp + i * 32) : 0ull;
  }
  // exclusive bin bases of this pass from the raw global histogram (every CTA redoes this 256-element scan)
  uint32_t gbase;
  {
    const uint32_t c = __ldg(ghist + tid);
    uint32_t inc = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t x = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += x;
    }
    if (lane == 31) s_wsum2[warp] = inc;
    __syncthreads();
    uint32_t pre = 0;
#pragma unroll
    for (int w = 0; w < WARPS; w++)
      if (w < warp) pre += s_wsum2[w];
    gbase = pre + inc - c;
  }
  // ---- digits (once), packed four to a register ----
  // padding items of a partial last tile take the highest digit and, being the last elements, rank last inside it
  uint32_t dig[ITEMS / 4];
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    uint32_t d = pass_digit(key[i], pd);
    if (!full && !(wofs + i * 32 < valid)) d = BINS - 1;
    if ((i & 3) == 0) dig[i >> 2] = d; else dig[i >> 2] |= d << (8 * (i & 3));
  }
  // ---- warp-level multi-split ranking (stable: items in ascending i, lanes in ascending order) ----
  uint16_t rank[ITEMS];
  uint32_t* wh = s_whist + warp * BINS;
  const uint32_t lt = (1u << lane) - 1;
#if LSR_SORT_RANK2
  // Two loops: the ballots of all items first (independent of one another, so they pipeline), then the running counts,
  // where only the lowest lane of a group of equal digits touches shared memory (load, add, store) and the others
  // receive its old value by shuffle — the chain from one item to the next is that lane's LDS -> STS, not
  // ballots -> LDS (all lanes) -> barrier -> STS -> barrier.
  uint32_t peers_of[ITEMS];
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    const uint32_t d = (dig[i >> 2] >> (8 * (i & 3))) & 0xffu;
    peers_of[i] = NB == 0 ? __match_any_sync(0xffffffffu, d) : (full ? match_digit<NB>(d) : match_digit<BITS>(d));
  }
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    const uint32_t d = (dig[i >> 2] >> (8 * (i & 3))) & 0xffu;
    const uint32_t peers = peers_of[i], lower = peers & lt;
    uint32_t prev = 0;
    if (lower == 0) {
      prev = wh[d];
      wh[d] = prev + __popc(peers);
    }
    __syncwarp();  // the store above is visible to whichever lane leads this digit's group in the next item
    prev = __shfl_sync(0xffffffffu, prev, __ffs(peers) - 1);
    rank[i] = (uint16_t)(prev + __popc(lower));
  }
#else
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    const uint32_t d = (dig[i >> 2] >> (8 * (i & 3))) & 0xffu;
    // (padding digits of a partial last tile may exceed NB bits, so that tile matches on all 8)
    const uint32_t peers = NB == 0 ? __match_any_sync(0xffffffffu, d) : (full ? match_digit<NB>(d) : match_digit<BITS>(d));
    const uint32_t lower = peers & lt;
    const uint32_t prev = wh[d];             // every lane of the group reads the running count ...
    __syncwarp();
    if (lower == 0) wh[d] = prev + __popc(peers);  // ... then its lowest lane bumps it
    rank[i] = (uint16_t)(prev + __popc(lower));
    __syncwarp();
  }
#endif
  __syncthreads();

  // ---- per bin (thread t == bin t): exclusive scan over warps, CTA bin scan, decoupled look-back ----
  uint32_t vcnt, bin_start;
  {
    const int t = tid;
    uint32_t wcnt[WARPS];
    uint32_t cnt = 0;
#pragma unroll
    for (int w = 0; w < WARPS; w++) {
      wcnt[w] = cnt;
      cnt += s_whist[w * BINS + t];
    }
    vcnt = cnt;
    if (t == BINS - 1) vcnt -= ((SORT_THREADS * ITEMS) - valid);
    volatile uint32_t* mine = lookback + (size_t)tile * BINS + t;
    *mine = (tile == 0 ? LB_INC : LB_AGG) | vcnt;  // publish before the CTA-wide scan
    uint32_t inc = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t x = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += x;
    }
    if (lane == 31) s_wsum[warp] = inc;
    __syncthreads();
    uint32_t pre = 0;
#pragma unroll
    for (int w = 0; w < WARPS; w++)
      if (w < warp) pre += s_wsum[w];
    bin_start = pre + inc - cnt;
    // s_whist[w][t] becomes the CTA-sorted position of warp w's first element of bin t
#pragma unroll
    for (int w = 0; w < WARPS; w++) s_whist[w * BINS + t] = bin_start + wcnt[w];
  }
  __syncthreads();

  // ---- stage keys in CTA-sorted order, write them out as coalesced runs ----
  // (rank[] now holds the CTA-sorted position, < (SORT_THREADS * ITEMS) <= 65536)
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    const uint32_t d = (dig[i >> 2] >> (8 * (i & 3))) & 0xffu;
    const uint32_t pos = wh[d] + rank[i];
    rank[i] = (uint16_t)pos;
    s_keys[sw(pos)] = key[i];
    s_dig[sw(pos)] = (uint8_t)d;
  }
  // the values are fetched now: their latency overlaps the key write-out
  uint32_t val[PAIRS ? ITEMS : 1];
  if (PAIRS) {
    const uint32_t* vp = vin + tile_base + wofs;
    if (full) {
#pragma unroll
      for (int i = 0; i < ITEMS; i++) val[i] = __ldg(vp + i * 32);
    } else {
#pragma unroll
      for (int i = 0; i < ITEMS; i++) val[i] = (wofs + i * 32 < valid) ? __ldg(vp + i * 32) : 0u;
    }
  }
  // ---- decoupled look-back (thread t == bin t); the staged keys and the value loads are already in flight ----
  {
    const int t = tid;
    volatile uint32_t* mine = lookback + (size_t)tile * BINS + t;
    uint32_t prefix = 0;
    if (tile != 0) {
      // look back LB_WIN tiles at a time (independent loads in flight), consume them nearest first.  (A window of 16
      // was measured, GPU call r2b: the R-sized passes 68 -> 73 us each, the single-wave depth passes unchanged —
      // they are bound by the ranking chain of a tile, not by the look-back.)
      int k = (int)tile - 1;
      bool done = false;
      while (!done) {
        uint32_t v[LB_WIN];
#pragma unroll
        for (int j = 0; j < LB_WIN; j++)
          v[j] = (k - j >= 0) ? lookback[(size_t)(k - j) * BINS + t] : (uint32_t)(2u << 30);
#pragma unroll
        for (int j = 0; j < LB_WIN; j++) {
          if (done) break;
          while ((v[j] & LB_FLAG) == 0) v[j] = lookback[(size_t)(k - j) * BINS + t];
          prefix += v[j] & LB_VAL;
          if ((v[j] & LB_FLAG) == LB_INC) done = true;
        }
        k -= LB_WIN;
      }
      *mine = LB_INC | (prefix + vcnt);
    }
    s_gofs[t] = gbase + prefix - bin_start;
  }

  __syncthreads();
  // Tile ranges straight out of the last pass of the instance sort (replaces identifyTileRanges,
  // rasterizer_impl.cu:116-138, and its extra sweep over the sorted list).  Neighbours in the CTA-sorted order that
  // share the digit are neighbours in the output, so a change of tile between them IS a range boundary: plain stores of
  // (start, ~end).  The first / last element of the CTA's chunk of a bin cannot see its neighbour in another CTA's
  // chunk; it offers its own position with atomicMin, and the true boundary — the smallest start, the smallest ~end —
  // wins whichever order the writes land in.  `ranges` is 0xFF-filled on entry and decoded by finalize_tiles_kernel.
  auto mark_range = [&](uint32_t p, uint64_t k, uint32_t dg, uint32_t g) {
    const uint32_t tile = (uint32_t)(k >> dyn.tile_shift);
    if (p == 0) {
      atomicMin(&dyn.ranges[tile].x, g);
    } else {
      const uint32_t ptile = (uint32_t)(s_keys[sw(p - 1)] >> dyn.tile_shift);
      const uint32_t pdg = s_dig[sw(p - 1)];
      if (pdg != dg) {
        atomicMin(&dyn.ranges[tile].x, g);
        atomicMin(&dyn.ranges[ptile].y, ~(s_gofs[pdg] + p));  // one past the predecessor's output position
      } else if (ptile != tile) {
        dyn.ranges[tile].x = g;
        dyn.ranges[ptile].y = ~g;
      }
    }
    if (p == valid - 1) atomicMin(&dyn.ranges[tile].y, ~(g + 1u));
  };
  const bool marking = !PAIRS && dyn.ranges != nullptr;
  if (full) {
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      const uint32_t p = tid + i * SORT_THREADS;
      const uint64_t k = s_keys[sw(p)];
      const uint32_t dg = s_dig[sw(p)];
      const uint32_t g = s_gofs[dg] + p;
      kout[g] = k;
      if (!PAIRS && ids_out != nullptr) ids_out[g] = (uint32_t)k & id_mask;
      if (marking) mark_range(p, k, dg, g);
    }
  } else {
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      const uint32_t p = tid + i * SORT_THREADS;
      if (p < valid) {
        const uint64_t k = s_keys[sw(p)];
        const uint32_t dg = s_dig[sw(p)];
        const uint32_t g = s_gofs[dg] + p;
        kout[g] = k;
        if (!PAIRS && ids_out != nullptr) ids_out[g] = (uint32_t)k & id_mask;
        if (marking) mark_range(p, k, dg, g);
      }
    }
  }
  if (!PAIRS) return;
  __syncthreads();
#pragma unroll
  for (int i = 0; i < ITEMS; i++) s_vals[sw(rank[i])] = val[PAIRS ? i : 0];
  __syncthreads();
  if (full) {
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      const uint32_t p = tid + i * SORT_THREADS;
      vout[s_gofs[s_dig[sw(p)]] + p] = s_vals[sw(p)];
    }
  } else {
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      const uint32_t p = tid + i * SORT_THREADS;
      if (p < valid) vout[s_gofs[s_dig[sw(p)]] + p] = s_vals[sw(p)];
    }
  }
}

using OnesweepFn = void (*)(const uint64_t*, uint64_t*, const uint32_t*, uint32_t*, uint32_t, uint32_t*,
                            volatile uint32_t*, const uint32_t*, PassDigit, uint32_t*, uint32_t, SortDynArgs);
template <bool PAIRS, int ITEMS>
OnesweepFn onesweep_for_bits_t(int nb) {
  switch (nb) {
    case 0: return onesweep_kernel<0, PAIRS, ITEMS>;
    case 1: return onesweep_kernel<1, PAIRS, ITEMS>;
    case 2: return onesweep_kernel<2, PAIRS, ITEMS>;
    case 3: return onesweep_kernel<3, PAIRS, ITEMS>;
    case 4: return onesweep_kernel<4, PAIRS, ITEMS>;
    case 5: return onesweep_kernel<5, PAIRS, ITEMS>;
    case 6: return onesweep_kernel<6, PAIRS, ITEMS>;
    case 7: return onesweep_kernel<7, PAIRS, ITEMS>;
    default: return onesweep_kernel<8, PAIRS, ITEMS>;
  }
}
// (1024-key tiles for small inputs were measured and rejected: the depth ordering of 1 M Gaussians went from 20 to
// 24 us per pass — the single-wave passes are bound by the look-back chain, which more tiles only lengthen)
OnesweepFn onesweep_for_bits(int nb, bool pairs, int items) {
  (void)items;
  return pairs ? onesweep_for_bits_t<true, SORT_ITEMS>(nb) : onesweep_for_bits_t<false, SORT_ITEMS>(nb);
}

}  // namespace

SortPlan make_sort_plan(uint32_t lo_min, uint32_t lo_max, int hi_bits, int id_bits) {
  SortPlan pl;
  pl.id_bits = id_bits;
  pl.lo_min = lo_min;
  uint32_t span = lo_max >= lo_min ? lo_max - lo_min : 0u;
  int lb = 0;
  while (lb < 32 && (span >> lb) != 0) lb++;
  pl.lo_bits = lb < 1 ? 1 : lb;
  pl.hi_bits = hi_bits < 0 ? 0 : hi_bits;
  pl.total_bits = pl.lo_bits + pl.hi_bits;
  if (pl.total_bits > 64) pl.total_bits = 64;
  pl.passes = (pl.total_bits + BITS - 1) / BITS;
  pl.packed = id_bits > 0 && pl.total_bits + id_bits <= 64;
  return pl;
}

SortPlan make_sort_plan_bits(int end_bit) {
  if (end_bit < 1) end_bit = 1;
  if (end_bit > 64) end_bit = 64;
  SortPlan pl;
  pl.lo_min = 0;
  pl.lo_bits = end_bit < 32 ? end_bit : 32;
  pl.hi_bits = end_bit - pl.lo_bits;
  pl.total_bits = end_bit;
  pl.passes = (end_bit + BITS - 1) / BITS;
  pl.id_bits = 0;
  pl.packed = false;
  return pl;
}

SortPlan make_sort_plan_packed(int key_bits, int id_bits) {
  if (key_bits < 1) key_bits = 1;
  SortPlan pl;
  pl.lo_min = 0;
  pl.lo_bits = 0;
  pl.hi_bits = key_bits;
  pl.total_bits = key_bits;
  pl.passes = (key_bits + BITS - 1) / BITS;
  pl.id_bits = id_bits;
  pl.packed = true;
  return pl;
}

void sort_prepare(void* work, int64_t n, const SortPlan& pl, cudaStream_t s) {
  // histograms + tickets + the look-back words of the passes that will run must be zero
  SortWork w = sort_work_view(work, n);
  const size_t lb_bytes = (size_t)pl.passes * sort_nblocks(n) * BINS * 4;
  cudaMemsetAsync(work, 0, (size_t)((char*)w.lookback - (char*)work) + lb_bytes, s);
}

int sort_pairs(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
               uint64_t* keys_tmp, uint32_t* vals_tmp, int64_t n, const SortPlan& pl, bool hist_ready, void* work,
               cudaStream_t s, const SortDyn* dyn) {
  if (n <= 0) return 0;
  SortWork w = sort_work_view(work, n);
  const int items = sort_items_for(n);
  const uint32_t nblocks = (uint32_t)((n + SORT_THREADS * items - 1) / (SORT_THREADS * items));
  static DeviceOnce attr;  // per device: a process may drive several GPUs
  attr.run([] {
    for (int nb = 0; nb <= BITS; nb++) {
      for (int it : {SORT_ITEMS}) {
        cudaFuncSetAttribute(onesweep_for_bits(nb, true, it), cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)sort_smem(it));
        cudaFuncSetAttribute(onesweep_for_bits(nb, false, it), cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)sort_smem(it));
      }
    }
  });
  const DigitMap dm{pl.lo_min, pl.lo_bits};
  if (!hist_ready) {
    sort_prepare(work, n, pl, s);
    const uint32_t hist_blocks = (uint32_t)std::min<int64_t>((n + 2047) / 2048, (int64_t)148 * 8);
    const size_t hist_smem = (size_t)8 * pl.passes * BINS * 4;
    cudaFuncSetAttribute(sort_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hist_smem);
    sort_hist_kernel<<<hist_blocks, 256, hist_smem, s>>>(keys_in, (uint32_t)n, pl.passes, pl.total_bits, dm,
                                                         pl.packed ? pl.id_bits : -1, w.hist);
    count_launch();
  }
  // Key width known on the device only (dyn->bits_dev, packed words): launch the worst-case number of passes; those
  // above the width return at once, the last one below it emits the ids.  The words then end up in keys_out or
  // keys_tmp depending on that width — such callers consume the ids only.
  const bool dyn_width = dyn != nullptr && dyn->bits_dev != nullptr;
  const int passes = dyn_width ? dyn->max_passes : pl.passes;
  const uint64_t* ksrc = keys_in;
  const uint32_t* vsrc = vals_in;
  for (int p = 0; p < passes; p++) {
    const bool to_out = dyn_width ? (p & 1) == 0 : ((passes - 1 - p) & 1) == 0;
    uint64_t* kdst = to_out ? keys_out : keys_tmp;
    uint32_t* vdst = to_out ? vals_out : vals_tmp;
    const int shift = p * BITS;
    const int bits = dyn_width ? BITS : std::min(BITS, pl.total_bits - shift);
    // d = (((lo - lo_min) >> s1) | (hi << s2) | (hi >> s3)) & mask of the compacted key (hi << lo_bits) | (lo - lo_min)
    PassDigit pd;
    pd.lo_min = pl.lo_min;
    pd.mask = (1u << bits) - 1;
    if (pl.packed) {
      // word = (compacted key << id_bits) | id:  d = (word >> (id_bits + shift)) & mask
      const int sh = pl.id_bits + shift;
      pd.lo_min = 0;
      pd.s1 = sh < 32 ? (uint32_t)sh : 32u;
      pd.s2 = (sh < 32 && sh > 0) ? (uint32_t)(32 - sh) : 32u;
      pd.s3 = sh >= 32 ? (uint32_t)(sh - 32) : 32u;
    } else {
      pd.s1 = shift < pl.lo_bits ? (uint32_t)shift : 32u;
      pd.s2 = (shift < pl.lo_bits && pl.lo_bits - shift < 32) ? (uint32_t)(pl.lo_bits - shift) : 32u;
      pd.s3 = shift >= pl.lo_bits ? (uint32_t)std::min(shift - pl.lo_bits, 32) : 32u;
    }
    const bool last = p == passes - 1;
    const bool depth_only = dyn_width || (shift + bits <= pl.lo_bits && bits == BITS);  // digit inside the depth word: MATCH.ANY
    SortDynArgs da{};
    da.shift = (uint32_t)shift;
    if (dyn != nullptr) {
      da.n_dev = dyn->n_dev;
      da.bits_dev = dyn->bits_dev;
      if (last && pl.packed) { da.ranges = dyn->ranges; da.tile_shift = dyn->tile_shift; }
    }
    uint32_t* ids = pl.packed && (last || dyn_width) ? vals_out : nullptr;
    onesweep_for_bits(depth_only ? LSR_DEPTH_NB : bits, !pl.packed, items)<<<nblocks, SORT_THREADS, sort_smem(items), s>>>(
        ksrc, kdst, vsrc, vdst, (uint32_t)n, w.tickets + p, w.lookback + (size_t)p * sort_nblocks(n) * BINS,
        w.hist + p * BINS, pd, ids, pl.id_bits >= 32 ? 0xffffffffu : ((1u << pl.id_bits) - 1), da);
    count_launch();
    ksrc = kdst;
    vsrc = vdst;
  }
  return passes;
}

// Allocation-free variant of the pair sort: same kernels, but the work area and the ping-pong buffers live in
// g_static_sort_scratch.  Every call shares that array (it keeps the last call's contents: sort_prepare re-zeroes the
// histograms, tickets and look-back words of the passes that will run; the ping-pong buffers are fully overwritten
// before they are read), so two calls that overlap on the GPU — different streams, or different host threads — race.
static cudaError_t static_sort_scratch(unsigned char** ptr) {
  // once per device, with that device current: the symbol has one address per device
  static std::mutex mu;
  static unsigned char* addr[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(mu);
  if (addr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_sort_scratch);
    if (e != cudaSuccess) return e;
    addr[dev] = static_cast<unsigned char*>(p);
  }
  *ptr = addr[dev];
  return cudaSuccess;
}

cudaError_t sort_pairs_static(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                              int64_t n, int end_bit, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  if (n > kStaticSortMaxN || sort_work_bytes(n) > kStaticWorkBytes) return cudaErrorInvalidValue;  // (the caller checks n)
  unsigned char* base = nullptr;
  cudaError_t e = static_sort_scratch(&base);
  if (e != cudaSuccess) return e;
  sort_pairs(keys_in, vals_in, keys_out, vals_out, reinterpret_cast<uint64_t*>(base + kStaticKeysOfs),
             reinterpret_cast<uint32_t*>(base + kStaticValsOfs), n, make_sort_plan_bits(end_bit), false, base, s);
  return cudaGetLastError();
}

}  // namespace lsr
