// binning.cu — K2' tile-count scan, K3' instance duplication, K5' tile ranges.
// Replaces cub::DeviceScan::InclusiveSum (RAST/cuda_rasterizer/rasterizer_impl.cu:280), duplicateWithKeys
// (:70-111) and identifyTileRanges (:116-138).  All HBM-bound integer work.
//
// Depth-first binning.  The reference sorts R (tile | depth) keys, one per Gaussian/tile instance (:306-311).  The
// same order is obtained with far less traffic by ordering the P Gaussians by depth first (a stable sort of P packed
// words, api.cu) and emitting the instances in that order: a stable sort of the instances on the TILE bits alone then
// leaves every tile's run in depth order, ties in Gaussian-index order — exactly the reference's stable result.
// The R-sized sort shrinks from ceil((tile_bits + depth_bits) / 8) passes to ceil(tile_bits / 8) (C2: 5 -> 2).
#include "common.cuh"
#include "gs_math.cuh"

namespace lsr {

namespace {

// One packed word per Gaussian for the depth ordering: ((depth bits - lo_min) << id_bits) | index.  Gaussians that
// touch no tile get depth field 0: where they land is irrelevant (they emit no instance).  The same pass counts the
// digits of every pass of the depth sort (shared-memory histograms, one flush per CTA).  The low digits of a float's
// bit pattern are close to uniform, so plain shared atomics rarely collide; the top digit takes few distinct values,
// so there the lanes holding the same digit are merged first (MATCH.ANY) and one of them adds the group's count.
constexpr int DW_THREADS = 256;
// The range of the depth word is only known on the device (ctrl, reduced by the preprocess kernel): every CTA derives
// lo_min and the width of the span from it, so the host never has to read it back; CTA 0 publishes the width for the
// sort passes (ctrl[CTRL_DEPTH_BITS]), which skip themselves when their digit lies above it.
__global__ void __launch_bounds__(DW_THREADS)
depth_words_kernel(int P, const float* __restrict__ depths, const uint32_t* __restrict__ tiles,
                   uint32_t* __restrict__ ctrl, int id_bits, uint64_t* __restrict__ words,
                   uint32_t* __restrict__ ghist) {
  __shared__ uint32_t s_hist[DEPTH_MAX_PASSES * SORT_BINS];
  const int tid = threadIdx.x, lane = tid & 31;
  const uint32_t lo_min = ~ctrl[CTRL_DEPTH_NEGMIN], lo_max = ctrl[CTRL_DEPTH_MAX];
  const uint32_t span = lo_max >= lo_min ? lo_max - lo_min : 0u;
  const int total_bits = max(1, 32 - __clz((int)span));
  const int passes = (total_bits + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS;
  if (blockIdx.x == 0 && tid == 0) ctrl[CTRL_DEPTH_BITS] = (uint32_t)total_bits;
  for (int j = tid; j < passes * SORT_BINS; j += DW_THREADS) s_hist[j] = 0;
  __syncthreads();
  // four elements per thread and iteration, all eight loads issued before the first use
  constexpr int DW_ITEMS = 4;
  const int stride = gridDim.x * DW_THREADS * DW_ITEMS;
  for (int base = blockIdx.x * DW_THREADS * DW_ITEMS; base < P; base += stride) {  // CTA-uniform trip count
    uint32_t t[DW_ITEMS], db[DW_ITEMS];
#pragma unroll
    for (int i = 0; i < DW_ITEMS; i++) {
      const int idx = base + i * DW_THREADS + tid;
      t[i] = idx < P ? __ldg(tiles + idx) : 0u;
      db[i] = idx < P ? __float_as_uint(__ldg(depths + idx)) : 0u;  // stale for culled Gaussians: masked by t == 0
    }
#pragma unroll
    for (int i = 0; i < DW_ITEMS; i++) {
      const int idx = base + i * DW_THREADS + tid;
      const bool ok = idx < P;
      const uint32_t d = t[i] != 0u ? db[i] - lo_min : 0u;
      if (ok) words[idx] = ((uint64_t)d << id_bits) | (uint64_t)(uint32_t)idx;
      for (int p = 0; p < passes; p++) {
        const int shift = p * SORT_RADIX_BITS;
        const uint32_t dg = (d >> shift) & (SORT_BINS - 1);  // bits above the span are zero
        if (p + 1 < passes) {
          if (ok) atomicAdd(&s_hist[p * SORT_BINS + dg], 1u);
        } else {
          const uint32_t peers = __match_any_sync(0xffffffffu, ok ? dg : 0xffffffffu);
          if (ok && (peers & ((1u << lane) - 1)) == 0)
            atomicAdd(&s_hist[p * SORT_BINS + dg], (uint32_t)__popc(peers));
        }
      }
    }
  }
  __syncthreads();
  for (int j = tid; j < passes * SORT_BINS; j += DW_THREADS)
    if (s_hist[j]) atomicAdd(&ghist[j], s_hist[j]);
}

// ------------------------------------------------------------------------------------------------------------
// K2': single-pass inclusive scan with decoupled look-back.  status word: [63:62] flag, [31:0] value.
// ------------------------------------------------------------------------------------------------------------
constexpr uint64_t FLAG_AGG = 1ull << 62, FLAG_INC = 2ull << 62, FLAG_MASK = 3ull << 62;

__global__ void __launch_bounds__(SCAN_THREADS)
scan_tiles_kernel(int P, const uint32_t* __restrict__ in, const uint32_t* __restrict__ perm,
                  uint32_t* __restrict__ out, volatile uint64_t* __restrict__ status, uint32_t* __restrict__ ctrl) {
  __shared__ uint32_t s_warp[SCAN_THREADS / 32];
  __shared__ uint32_t s_bid, s_prefix;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_bid = atomicAdd(&ctrl[CTRL_SCAN_TICKET], 1u);
  __syncthreads();
  const uint32_t bid = s_bid;
  const int base = bid * SCAN_TILE + tid * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS];
  if (perm != nullptr) {  // scan in[perm[i]] (the depth order)
    if (base + SCAN_ITEMS <= P) {
      const uint4 a = __ldg(reinterpret_cast<const uint4*>(perm + base));
      const uint4 b = __ldg(reinterpret_cast<const uint4*>(perm + base) + 1);
      v[0] = __ldg(in + a.x); v[1] = __ldg(in + a.y); v[2] = __ldg(in + a.z); v[3] = __ldg(in + a.w);
      v[4] = __ldg(in + b.x); v[5] = __ldg(in + b.y); v[6] = __ldg(in + b.z); v[7] = __ldg(in + b.w);
    } else {
#pragma unroll
      for (int i = 0; i < SCAN_ITEMS; i++) v[i] = (base + i < P) ? in[perm[base + i]] : 0u;
    }
  } else if (base + SCAN_ITEMS <= P) {
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(in + base));
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(in + base) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) v[i] = (base + i < P) ? in[base + i] : 0u;
  }
  uint32_t tsum = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; i++) { tsum += v[i]; v[i] = tsum; }
  // warp inclusive scan of thread sums
  uint32_t w = tsum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t n = __shfl_up_sync(0xffffffffu, w, d);
    if (lane >= d) w += n;
  }
  if (lane == 31) s_warp[warp] = w;
  __syncthreads();
  uint32_t warp_prefix = 0, block_sum = 0;
#pragma unroll
  for (int i = 0; i < SCAN_THREADS / 32; i++) {
    const uint32_t x = s_warp[i];
    if (i < warp) warp_prefix += x;
    block_sum += x;
  }
  // decoupled look-back by warp 0
  if (warp == 0) {
    if (lane == 0) {
      status[bid] = (bid == 0 ? FLAG_INC : FLAG_AGG) | block_sum;
    }
    uint32_t prefix = 0;
    if (bid > 0) {
      int look = (int)bid - 1;
      while (true) {
        const int j = look - lane;
        uint64_t st = 0;
        if (j >= 0) {
          do { st = status[j]; } while ((st & FLAG_MASK) == 0);
        } else {
          st = FLAG_INC;  // virtual zero-prefix before block 0
        }
        const uint32_t inc_mask = __ballot_sync(0xffffffffu, (st & FLAG_MASK) == FLAG_INC);
        // nearest predecessor with an inclusive prefix; none in this window -> take all 32 aggregates
        const int first_inc = inc_mask ? __ffs(inc_mask) - 1 : 31;
        uint32_t val = (lane <= first_inc) ? (uint32_t)st : 0u;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) val += __shfl_xor_sync(0xffffffffu, val, d);
        prefix += val;
        if (inc_mask) break;
        look -= 32;
      }
      if (lane == 0) {
        status[bid] = FLAG_INC | (uint64_t)(prefix + block_sum);
      }
    }
    if (lane == 0) {
      s_prefix = prefix;
      if (perm == nullptr && (int)(bid + 1) * SCAN_TILE >= P) ctrl[CTRL_NUM_RENDERED] = prefix + block_sum;
    }
  }
  __syncthreads();
  const uint32_t add = s_prefix + warp_prefix + (w - tsum);
  if (base + SCAN_ITEMS <= P) {
    uint4 a = make_uint4(v[0] + add, v[1] + add, v[2] + add, v[3] + add);
    uint4 b = make_uint4(v[4] + add, v[5] + add, v[6] + add, v[7] + add);
    reinterpret_cast<uint4*>(out + base)[0] = a;
    reinterpret_cast<uint4*>(out + base)[1] = b;
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++)
      if (base + i < P) out[base + i] = v[i] + add;
  }
}

// ------------------------------------------------------------------------------------------------------------
// K3': one word per Gaussian/tile overlap, word = (tile_id << id_bits) | Gaussian index; rect walked y-major then x
// (rasterizer_impl.cu:85-109), Gaussians taken in depth order (perm), instances of Gaussian perm[i] written
// consecutively from offsets[i-1].  The reference's key tile_id << 32 | float_bits(depth) is implied: the depth
// ordering is already in the emission order, and the depth bits are depths[index] (lsr_debug_sorted_keys).
//
// Load-balanced: a warp owns 32 consecutive Gaussians (of the depth order), whose instances occupy one contiguous
// output range; the 32 lanes walk that range together (lane = consecutive output slot), finding the owning Gaussian
// of each slot with a 5-step binary search over the lanes' start offsets (shuffles).  Every store instruction
// therefore writes 32 consecutive words (256 B) — the reference's one-thread-per-Gaussian loop scatters them.
// The division by the rect width uses a per-Gaussian magic reciprocal (exact for slot, width < 2^16).
//
// The kernel also accumulates the tile sort's digit histograms for every pass (in shared memory, flushed once per
// CTA), so the sort never reads the words just to count them; runs of equal digits in a warp are collapsed into one
// shared-memory atomic (a full-warp shared atomic costs ~34 SM-cycles on B200, so only run heads issue one).
// ------------------------------------------------------------------------------------------------------------
constexpr int DUP_THREADS = 256;

__global__ void __launch_bounds__(DUP_THREADS)
duplicate_kernel(int P, int gx, int gy, const uint32_t* __restrict__ perm, const int32_t* __restrict__ radii,
                 const GeomRec* __restrict__ rec, const uint32_t* __restrict__ offsets, uint64_t* __restrict__ words,
                 uint32_t capacity, SortPlan plan, uint32_t* __restrict__ ghist) {
  __shared__ uint32_t s_hist[SORT_MAX_PASSES * SORT_BINS];
  const int tid = threadIdx.x, lane = tid & 31;
  const int passes = plan.passes;
  for (int j = tid; j < passes * SORT_BINS; j += DUP_THREADS) s_hist[j] = 0;
  __syncthreads();
  const int nblocks = (P + DUP_THREADS - 1) / DUP_THREADS;
  for (int blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
    const int i = blk * DUP_THREADS + tid;
    uint32_t x0 = 0, y0 = 0, x1 = 0, y1 = 0, incl = 0, gid = 0;
    int radius = 0;
    if (i < P) {
      gid = perm[i];
      radius = radii[gid];
    }
    incl = offsets[min(i, P - 1)];  // lanes past the end sit at the end of the output range with zero instances
    if (radius > 0) {
      const float2 xy = *reinterpret_cast<const float2*>(rec + gid);
      get_rect(xy.x, xy.y, radius, gx, gy, x0, y0, x1, y1);
    }
    const uint32_t w = x1 - x0;
    const uint32_t cnt = w * (y1 - y0);
    const uint32_t start = incl - cnt;                                  // exclusive offset of this Gaussian
    const uint32_t wstart = __shfl_sync(0xffffffffu, start, 0);          // warp's output range [wstart, wend)
    const uint32_t wend = __shfl_sync(0xffffffffu, incl, 31);
    const uint32_t rel = start - wstart;
    const uint32_t magic = w > 1 ? 0xffffffffu / w + 1u : 0u;  // ceil(2^32 / w)
    const uint32_t origin = (y0 << 16) | x0;
    const uint32_t total = wend - wstart;
    for (uint32_t k0 = 0; k0 < total; k0 += 32) {
      const uint32_t k = k0 + lane;
      // last lane j with rel_j <= k (zero-count lanes share their successor's start, so they are never picked)
      int j = 0;
#pragma unroll
      for (int step = 16; step > 0; step >>= 1) {
        const uint32_t r = __shfl_sync(0xffffffffu, rel, (j + step) & 31);
        if (r <= k) j += step;
      }
      const uint32_t jrel = __shfl_sync(0xffffffffu, rel, j);
      const uint32_t jw = __shfl_sync(0xffffffffu, w, j);
      const uint32_t jmagic = __shfl_sync(0xffffffffu, magic, j);
      const uint32_t jorigin = __shfl_sync(0xffffffffu, origin, j);
      const uint32_t jgid = __shfl_sync(0xffffffffu, gid, j);
      const bool ok = k < total;
      const uint32_t r = k - jrel;
      const uint32_t q = jw > 1 ? __umulhi(r, jmagic) : r;  // r / jw
      const uint32_t y = (jorigin >> 16) + q, x = (jorigin & 0xffffu) + (r - q * jw);
      const uint32_t tile = y * (uint32_t)gx + x;
      const uint32_t o = wstart + k;
      if (ok && o < capacity) words[o] = ((uint64_t)tile << plan.id_bits) | (uint64_t)jgid;
      // digits of the tile id: one shared atomic per run of equal digits
      for (int p = 0; p < passes; p++) {
        const int shift = p * SORT_RADIX_BITS;
        const int bits = min(SORT_RADIX_BITS, plan.total_bits - shift);
        const uint32_t d = ok ? ((tile >> shift) & ((1u << bits) - 1)) : 0xffffffffu;
        const uint32_t prev = __shfl_up_sync(0xffffffffu, d, 1);
        const bool head = (lane == 0) || (prev != d);
        const uint32_t heads = __ballot_sync(0xffffffffu, head);
        if (head && ok) {
          const uint32_t later = lane == 31 ? 0u : (heads >> (lane + 1));
          const int run = later ? __ffs(later) : (32 - lane);
          atomicAdd(&s_hist[p * SORT_BINS + d], (uint32_t)run);
        }
      }
    }
  }
  __syncthreads();
  for (int j = tid; j < passes * SORT_BINS; j += DUP_THREADS)
    if (s_hist[j]) atomicAdd(&ghist[j], s_hist[j]);
}

// ------------------------------------------------------------------------------------------------------------
// K2'+K3' fused: duplication that computes its own offsets.  A block takes 256 consecutive Gaussians of the depth
// order (dynamic ticket, so blocks start in order), scans their tile counts, obtains its base from the blocks before
// it by decoupled look-back (one warp, 32 predecessors per poll), and emits its instances: no scan kernel, no
// offsets round trip through HBM, no binary search per slot.
//
// Emission goes through shared memory in chunks of DUP_CHUNK slots: every thread writes the words of ITS Gaussian that
// fall into the chunk (row-major walk of the rect with two counters: no division in the loop), then the block copies
// the chunk out with fully coalesced 8-byte stores and counts the tile sort's digits on the way.  Gaussians with more
// than DUP_BIG instances would make one lane loop alone, so the block emits those cooperatively (slot -> row, column
// by one division per slot).
//
// Nothing here needs the instance count on the host: stores are clipped at `capacity`, and an overflow raises
// ERR_OVERFLOW in ctrl (the forward then poisons its outputs, api.cu).
//
// Measured and dropped (GPU calls r2j, r2k; ncu on this version shows 12 barrier-stall cycles per issued instruction,
// the block waiting for warp 0's look-back): filling the first chunk beside the look-back with 3 072-slot chunks and
// half the barriers: 97 us instead of 87 at C2, 602 instead of 501 at C5; additionally taking the tickets one block
// ahead to prefetch the (index, radius, centre) gathers: 112 / 1 214 us — a block that is ticketed but idle holds up
// the look-back of every block behind it.
// ------------------------------------------------------------------------------------------------------------
constexpr int DUP_CHUNK = 8 * DUP_BLOCK;  // staged words per chunk (16 KB)
constexpr uint32_t DUP_BIG = 512;

__global__ void __launch_bounds__(DUP_BLOCK)
dup_scan_kernel(int P, int gx, int gy, const uint32_t* __restrict__ perm, const int32_t* __restrict__ radii,
                const GeomRec* __restrict__ rec, uint32_t* __restrict__ offsets, uint64_t* __restrict__ words,
                uint32_t capacity, int id_bits, int passes, volatile uint64_t* __restrict__ status,
                uint32_t* __restrict__ ctrl, uint32_t* __restrict__ ghist) {
  __shared__ __align__(16) uint64_t s_words[DUP_CHUNK];
  __shared__ uint32_t s_hist[SORT_MAX_PASSES * SORT_BINS];
  __shared__ uint32_t s_warp[DUP_BLOCK / 32];
  __shared__ uint32_t s_blk, s_base, s_nbig;
  __shared__ uint4 s_bigrec[DUP_BLOCK];  // Gaussians emitted cooperatively: block-local start, count, width, y0<<16|x0
  __shared__ uint32_t s_biggid[DUP_BLOCK];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int j = tid; j < passes * SORT_BINS; j += DUP_BLOCK) s_hist[j] = 0;
  if (blockIdx.x == 0 && tid == 0 && ctrl[CTRL_NUM_RENDERED] > capacity) atomicOr(&ctrl[CTRL_ERROR], ERR_OVERFLOW);
  const int nblocks = (P + DUP_BLOCK - 1) / DUP_BLOCK;
  for (;;) {
    __syncthreads();  // previous block's shared state is no longer read
    if (tid == 0) { s_blk = atomicAdd(&ctrl[CTRL_DUP_TICKET], 1u); s_nbig = 0; }
    __syncthreads();
    const uint32_t blk = s_blk;
    if (blk >= (uint32_t)nblocks) break;
    const int i = (int)blk * DUP_BLOCK + tid;
    uint32_t x0 = 0, y0 = 0, x1 = 0, y1 = 0, gid = 0;
    int radius = 0;
    if (i < P) {
      gid = __ldg(perm + i);
      radius = __ldg(radii + gid);
    }
    if (radius > 0) {
      const float2 xy = __ldg(reinterpret_cast<const float2*>(rec + gid));
      get_rect(xy.x, xy.y, radius, gx, gy, x0, y0, x1, y1);
    }
    const uint32_t w = x1 - x0;
    const uint32_t cnt = w * (y1 - y0);  // == tiles_touched of this Gaussian
    // ---- block scan of the counts ----
    uint32_t inc = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t n = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += n;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t warp_prefix = 0, block_sum = 0;
#pragma unroll
    for (int k = 0; k < DUP_BLOCK / 32; k++) {
      const uint32_t x = s_warp[k];
      if (k < warp) warp_prefix += x;
      block_sum += x;
    }
    // ---- base of the block: decoupled look-back by warp 0 (status word: [63:62] flag, [31:0] value) ----
    if (warp == 0) {
      if (lane == 0) status[blk] = (blk == 0 ? FLAG_INC : FLAG_AGG) | block_sum;
      uint32_t prefix = 0;
      if (blk > 0) {
        int look = (int)blk - 1;
        while (true) {
          const int j = look - lane;
          uint64_t st = FLAG_INC;  // virtual zero prefix before block 0
          if (j >= 0) {
            do { st = status[j]; } while ((st & FLAG_MASK) == 0);
          }
          const uint32_t inc_mask = __ballot_sync(0xffffffffu, (st & FLAG_MASK) == FLAG_INC);
          const int first_inc = inc_mask ? __ffs(inc_mask) - 1 : 31;
          uint32_t val = (lane <= first_inc) ? (uint32_t)st : 0u;
#pragma unroll
          for (int d = 16; d > 0; d >>= 1) val += __shfl_xor_sync(0xffffffffu, val, d);
          prefix += val;
          if (inc_mask) break;
          look -= 32;
        }
        if (lane == 0) status[blk] = FLAG_INC | (uint64_t)(prefix + block_sum);
      }
      if (lane == 0) s_base = prefix;
    }
    const uint32_t lend = warp_prefix + inc;  // block-local [lstart, lend) of this thread's instances
    const uint32_t lstart = lend - cnt;
    const uint32_t origin = (y0 << 16) | x0;
    if (cnt > DUP_BIG) {
      const uint32_t k = atomicAdd(&s_nbig, 1u);
      s_bigrec[k] = make_uint4(lstart, cnt, w, origin);
      s_biggid[k] = gid;
    }
    __syncthreads();
    const uint32_t base = s_base, nbig = s_nbig;
    if (i < P) offsets[i] = base + lend;  // inclusive, as the stand-alone scan wrote it
    // ---- emission, DUP_CHUNK slots at a time ----
    for (uint32_t c0 = 0; c0 < block_sum; c0 += DUP_CHUNK) {
      const uint32_t c1 = min(c0 + (uint32_t)DUP_CHUNK, block_sum);
      if (cnt != 0u && cnt <= DUP_BIG && lstart < c1 && lend > c0) {
        const uint32_t a = max(lstart, c0), b = min(lend, c1);
        uint32_t r = a - lstart;
        uint32_t row = 0, col = r;
        if (r != 0u) { row = r / w; col = r - row * w; }  // only for a Gaussian cut by the chunk boundary
        uint32_t tile = (y0 + row) * (uint32_t)gx + x0 + col;
        for (uint32_t sl = a; sl < b; sl++) {
          s_words[sl - c0] = ((uint64_t)tile << id_bits) | (uint64_t)gid;
          col++; tile++;
          if (col == w) { col = 0; tile += (uint32_t)gx - w; }
        }
      }
      for (uint32_t bi = 0; bi < nbig; bi++) {  // the block's big Gaussians, all threads together
        const uint4 br = s_bigrec[bi];
        const uint32_t a = max(br.x, c0), b = min(br.x + br.y, c1);
        const uint32_t bgid = s_biggid[bi];
        for (uint32_t sl = a + tid; sl < b; sl += DUP_BLOCK) {
          const uint32_t r = sl - br.x, row = r / br.z, col = r - row * br.z;
          const uint32_t tile = ((br.w >> 16) + row) * (uint32_t)gx + (br.w & 0xffffu) + col;
          s_words[sl - c0] = ((uint64_t)tile << id_bits) | (uint64_t)bgid;
        }
      }
      __syncthreads();
      const uint32_t n = c1 - c0;
      for (uint32_t k0 = 0; k0 < n; k0 += DUP_BLOCK) {  // CTA-uniform trip count (warp-wide votes below)
        const uint32_t k = k0 + tid;
        const uint32_t o = base + c0 + k;
        const bool ok = k < n && o < capacity;  // clipped instances are neither stored nor counted
        const uint64_t word = ok ? s_words[k] : 0ull;
        if (ok) words[o] = word;
        const uint32_t tile = (uint32_t)(word >> id_bits);
        // digits of the tile id for every pass of the tile sort.  Lowest digit: consecutive slots hold consecutive
        // tiles, so plain shared atomics rarely collide.  Higher digits: a warp's slots mostly share one value, so the
        // lanes holding the same digit are merged first and one of them adds the group's count.
        if (ok) atomicAdd(&s_hist[tile & (SORT_BINS - 1)], 1u);
        for (int p = 1; p < passes; p++) {
          const uint32_t d = ok ? ((tile >> (p * SORT_RADIX_BITS)) & (SORT_BINS - 1)) : 0xffffffffu;
          const uint32_t peers = __match_any_sync(0xffffffffu, d);
          if (ok && (peers & ((1u << lane) - 1)) == 0) atomicAdd(&s_hist[p * SORT_BINS + d], (uint32_t)__popc(peers));
        }
      }
      __syncthreads();
    }
  }
  // (all threads left the loop together after the barrier that published s_blk)
  for (int j = tid; j < passes * SORT_BINS; j += DUP_BLOCK)
    if (s_hist[j]) atomicAdd(&ghist[j], s_hist[j]);
}

// ------------------------------------------------------------------------------------------------------------
// K5': identifyTileRanges (rasterizer_impl.cu:116-138).  ranges must be zero on entry (cudaMemset, :313).
// ------------------------------------------------------------------------------------------------------------
// The blend kernels' CTA -> tile map: tiles by descending list length, so that the long-running quadrants start first
// and the last wave of the grid consists of short ones (the hardware hands out CTAs in index order).  Counting sort on
// a 256-level logarithmic length scale by ONE CTA of 1024 threads; the order inside a bucket is irrelevant (it only schedules; results do
// not depend on it).
// 256 buckets on a logarithmic scale (4 fractional bits: ~6 % resolution at any list length), longest lists first
__device__ __forceinline__ uint32_t order_bucket(uint2 r) {
  const uint32_t len = r.y - r.x;
  uint32_t key = len;
  if (len >= 16u) {
    const uint32_t e = 31u - (uint32_t)__clz((int)len);  // >= 4
    key = min((e - 3u) * 16u + ((len >> (e - 4u)) & 15u), 255u);
  }
  return 255u - key;
}

// ENC: the ranges come straight out of the last sort pass (radix_sort.cu), which marks them with min-operations on
// (start, ~end) over a 0xFF-filled array; they are decoded in place here — (0, 0) for a tile nobody touched, as the
// reference's zero-initialised ranges read (rasterizer_impl.cu:313) — by the thread that then buckets them.
template <bool ENC>
__device__ __forceinline__ void order_tiles(int T, uint2* __restrict__ ranges, uint32_t* __restrict__ order,
                                            uint32_t* s_cnt) {
  const int tid = threadIdx.x;
  if (tid < 256) s_cnt[tid] = 0;
  __syncthreads();
  constexpr int CH = 8;  // tiles per thread and chunk: the 8 loads of a chunk are in flight together
  for (int base = 0; base < T; base += CH * 1024) {
    uint32_t bucket[CH];
#pragma unroll
    for (int k = 0; k < CH; k++) {
      const int t = base + tid + k * 1024;
      bucket[k] = 256u;
      if (t < T) {
        uint2 r = ranges[t];
        if (ENC) {
          r = r.x == 0xffffffffu ? make_uint2(0u, 0u) : make_uint2(r.x, ~r.y);
          ranges[t] = r;
        }
        bucket[k] = order_bucket(r);
      }
    }
#pragma unroll
    for (int k = 0; k < CH; k++)  // (warp-aggregating these with MATCH.ANY was measured slower)
      if (bucket[k] < 256u) atomicAdd(&s_cnt[bucket[k]], 1u);
  }
  __syncthreads();
  if (tid < 32) {  // exclusive scan of the 256 counts by one warp (8 per lane)
    uint32_t v[8], sum = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) { v[i] = s_cnt[tid * 8 + i]; sum += v[i]; }
    uint32_t inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
      if (tid >= d) inc += o;
    }
    uint32_t run = inc - sum;
#pragma unroll
    for (int i = 0; i < 8; i++) { s_cnt[tid * 8 + i] = run; run += v[i]; }
  }
  __syncthreads();
  for (int base = 0; base < T; base += CH * 1024) {
    uint32_t bucket[CH];
#pragma unroll
    for (int k = 0; k < CH; k++) {
      const int t = base + tid + k * 1024;
      bucket[k] = t < T ? order_bucket(ranges[t]) : 256u;  // (written above by this same thread when ENC)
    }
#pragma unroll
    for (int k = 0; k < CH; k++)
      if (bucket[k] < 256u) order[atomicAdd(&s_cnt[bucket[k]], 1u)] = (uint32_t)(base + tid + k * 1024);
  }
}

__global__ void __launch_bounds__(1024)
tile_order_kernel(int T, uint2* __restrict__ ranges, uint32_t* __restrict__ order) {
  __shared__ uint32_t s_cnt[256];
  order_tiles<false>(T, ranges, order, s_cnt);
}

// Last binning step of a forward: decode the ranges, order the tiles, and keep what the device found out about this
// frame (instance count, error flags, depth range) with the call's own state, where lsr_forward_status / the
// asynchronous host copy pick it up.
__global__ void __launch_bounds__(1024)
finalize_tiles_kernel(int T, uint2* __restrict__ ranges, uint32_t* __restrict__ order, bool encoded,
                      const uint32_t* __restrict__ ctrl, uint32_t capacity, uint32_t stamp,
                      uint32_t* __restrict__ status) {
  __shared__ uint32_t s_cnt[256];
  if (threadIdx.x == 0) {
    status[LSR_STATUS_NUM_RENDERED] = ctrl[CTRL_NUM_RENDERED];
    status[LSR_STATUS_FLAGS] = ctrl[CTRL_ERROR];
    status[LSR_STATUS_DEPTH_MIN] = ~ctrl[CTRL_DEPTH_NEGMIN];
    status[LSR_STATUS_DEPTH_MAX] = ctrl[CTRL_DEPTH_MAX];
    status[LSR_STATUS_CAPACITY] = capacity;
    status[LSR_STATUS_STAMP] = stamp;
  }
  if (encoded) order_tiles<true>(T, ranges, order, s_cnt);
  else order_tiles<false>(T, ranges, order, s_cnt);
}

// An instance list that did not fit leaves incomplete images behind: overwrite them with NaN so that they can never
// pass for a render (asynchronous forward only; the synchronous one returns LSR_ERR_CAPACITY before blending).
__global__ void __launch_bounds__(256)
poison_outputs_kernel(const uint32_t* __restrict__ status, float* __restrict__ a, size_t na, float* __restrict__ b,
                      size_t nb) {
  if ((status[LSR_STATUS_FLAGS] & ERR_OVERFLOW) == 0u) return;
  const float nan = __int_as_float(0x7fc00000);
  const size_t stride = (size_t)gridDim.x * 256;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < na; i += stride) a[i] = nan;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < nb; i += stride) b[i] = nan;
}

// Four consecutive words per thread (two 128-bit loads), one extra load for the predecessor of the first.
__device__ __forceinline__ void mark_ranges(uint32_t i0, uint32_t L, const uint64_t* __restrict__ keys, int tile_shift,
                                            uint2* __restrict__ ranges) {
  uint64_t k[4];
  if (i0 + 4 <= L) {
    const ulonglong2 a = __ldg(reinterpret_cast<const ulonglong2*>(keys + i0));
    const ulonglong2 b = __ldg(reinterpret_cast<const ulonglong2*>(keys + i0) + 1);
    k[0] = a.x; k[1] = a.y; k[2] = b.x; k[3] = b.y;
  } else {
#pragma unroll
    for (int j = 0; j < 4; j++) k[j] = i0 + j < L ? keys[i0 + j] : 0ull;
  }
  uint32_t prev = i0 == 0 ? 0xffffffffu : (uint32_t)(__ldg(keys + i0 - 1) >> tile_shift);
#pragma unroll
  for (int j = 0; j < 4; j++) {
    const uint32_t idx = i0 + j;
    if (idx >= L) break;
    const uint32_t cur = (uint32_t)(k[j] >> tile_shift);
    if (cur != prev) {
      if (idx != 0) ranges[prev].y = idx;
      ranges[cur].x = idx;
    }
    if (idx == L - 1) ranges[cur].y = L;
    prev = cur;
  }
}

// (Ordering the tiles inside this kernel — the CTA that finishes last does it — was measured: the ticket and the fence
// cost every CTA more than the extra launch saves, 32.6 vs 25.5 us for the stage.)
__global__ void __launch_bounds__(256)
tile_ranges_kernel(uint32_t L, const uint64_t* __restrict__ keys, int tile_shift, uint2* __restrict__ ranges) {
  const uint32_t i0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4u;
  if (i0 < L) mark_ranges(i0, L, keys, tile_shift, ranges);
}

// sorted word (tile << id_bits | index) -> the reference's key format (tile << 32 | depth bits of that Gaussian)
__global__ void __launch_bounds__(256)
unpack_keys_kernel(uint32_t L, const uint64_t* __restrict__ words, const float* __restrict__ depths,
                   uint64_t* __restrict__ keys, int id_bits) {
  const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= L) return;
  const uint64_t w = words[idx];
  const uint32_t gid = (uint32_t)(w & ((1ull << id_bits) - 1));
  keys[idx] = ((w >> id_bits) << 32) | (uint64_t)__float_as_uint(depths[gid]);
}

}  // namespace

void launch_depth_words(int P, const float* depths, const uint32_t* tiles, uint32_t* ctrl, int id_bits, uint64_t* words,
                        uint32_t* ghist, cudaStream_t s) {
  if (P <= 0) return;
  const int nblocks = (P + DW_THREADS * 4 - 1) / (DW_THREADS * 4);
  const int grid = nblocks < 148 * 4 ? nblocks : 148 * 4;  // persistent CTAs: each flushes passes*256 global atomics
  depth_words_kernel<<<grid, DW_THREADS, 0, s>>>(P, depths, tiles, ctrl, id_bits, words, ghist);
  count_launch();
}

void launch_dup_scan(int P, int grid_x, int grid_y, const uint32_t* perm, const int32_t* radii, GeomView g,
                     uint64_t* words, int64_t capacity, const SortPlan& plan, uint64_t* status, uint32_t* ctrl,
                     uint32_t* ghist, cudaStream_t s) {
  if (P <= 0) return;
  const int nblocks = (P + DUP_BLOCK - 1) / DUP_BLOCK;
  const int per_sm = 6 * 256 / DUP_BLOCK;
  const int grid = nblocks < 148 * per_sm ? nblocks : 148 * per_sm;  // persistent CTAs (29 KB of shared memory each)
  dup_scan_kernel<<<grid, DUP_BLOCK, 0, s>>>(P, grid_x, grid_y, perm, radii, g.rec, g.offsets, words,
                                               (uint32_t)capacity, plan.id_bits, plan.passes, status, ctrl, ghist);
  count_launch();
}

void launch_scan_tiles(int P, const uint32_t* tiles, const uint32_t* perm, uint32_t* offsets, uint64_t* status,
                       uint32_t* ctrl, cudaStream_t s) {
  if (P <= 0) return;
  const int blocks = (P + SCAN_TILE - 1) / SCAN_TILE;
  scan_tiles_kernel<<<blocks, SCAN_THREADS, 0, s>>>(P, tiles, perm, offsets, status, ctrl);
  count_launch();
}

void launch_duplicate(int P, int grid_x, int grid_y, const uint32_t* perm, const int32_t* radii, GeomView g,
                      uint64_t* words, int64_t capacity, const SortPlan& plan, uint32_t* ghist, cudaStream_t s) {
  if (P <= 0) return;
  const int nblocks = (P + DUP_THREADS - 1) / DUP_THREADS;
  const int grid = nblocks < 148 * 8 ? nblocks : 148 * 8;  // persistent CTAs: one histogram flush each
  duplicate_kernel<<<grid, DUP_THREADS, 0, s>>>(P, grid_x, grid_y, perm, radii, g.rec, g.offsets, words,
                                                   (uint32_t)capacity, plan, ghist);
  count_launch();
}

void launch_tile_ranges(int64_t R, const uint64_t* keys, int tile_shift, uint2* ranges, cudaStream_t s) {
  if (R <= 0) return;
  tile_ranges_kernel<<<(unsigned)((R + 1023) / 1024), 256, 0, s>>>((uint32_t)R, keys, tile_shift, ranges);
  count_launch();
}

void launch_tile_order(int T, uint2* ranges, uint32_t* order, cudaStream_t s) {
  if (T <= 0) return;
  tile_order_kernel<<<1, 1024, 0, s>>>(T, ranges, order);
  count_launch();
}

void launch_finalize_tiles(int T, uint2* ranges, uint32_t* order, bool encoded, const uint32_t* ctrl,
                           int64_t capacity, uint32_t stamp, uint32_t* status, cudaStream_t s) {
  finalize_tiles_kernel<<<1, 1024, 0, s>>>(T, ranges, order, encoded, ctrl, (uint32_t)capacity, stamp, status);
  count_launch();
}

void launch_poison_outputs(const uint32_t* status, float* a, size_t na, float* b, size_t nb, cudaStream_t s) {
  poison_outputs_kernel<<<148 * 2, 256, 0, s>>>(status, a, a ? na : 0, b, b ? nb : 0);
  count_launch();
}

void launch_unpack_keys(int64_t R, const uint64_t* words, const float* depths, uint64_t* keys, int id_bits,
                        cudaStream_t s) {
  if (R <= 0) return;
  unpack_keys_kernel<<<(unsigned)((R + 255) / 256), 256, 0, s>>>((uint32_t)R, words, depths, keys, id_bits);
  count_launch();
}

}  // namespace lsr
