// binning.cu — K2' tile-count scan, K3' key/value duplication, K5' tile ranges.
// Replaces cub::DeviceScan::InclusiveSum (RAST/cuda_rasterizer/rasterizer_impl.cu:280), duplicateWithKeys
// (:70-111) and identifyTileRanges (:116-138).  All HBM-bound integer work.
#include "common.cuh"
#include "gs_math.cuh"

namespace lsr {

namespace {

// ------------------------------------------------------------------------------------------------------------
// K2': single-pass inclusive scan with decoupled look-back.  status word: [63:62] flag, [31:0] value.
// ------------------------------------------------------------------------------------------------------------
constexpr uint64_t FLAG_AGG = 1ull << 62, FLAG_INC = 2ull << 62, FLAG_MASK = 3ull << 62;

__global__ void __launch_bounds__(SCAN_THREADS)
scan_tiles_kernel(int P, const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                  volatile uint64_t* __restrict__ status, uint32_t* __restrict__ ctrl) {
  __shared__ uint32_t s_warp[SCAN_THREADS / 32];
  __shared__ uint32_t s_bid, s_prefix;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_bid = atomicAdd(&ctrl[CTRL_SCAN_TICKET], 1u);
  __syncthreads();
  const uint32_t bid = s_bid;
  const int base = bid * SCAN_TILE + tid * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS];
  if (base + SCAN_ITEMS <= P) {
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
      if ((int)(bid + 1) * SCAN_TILE >= P) ctrl[CTRL_NUM_RENDERED] = prefix + block_sum;
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
// K3': one (key, value) pair per Gaussian/tile overlap; key = tile_id << 32 | float_bits(depth); rect walked y-major
// then x, pairs written consecutively from offsets[idx-1]  (rasterizer_impl.cu:85-109).
// Warp-cooperative: the 32 Gaussians of a warp are emitted one after another by all 32 lanes, so a large rect is
// written with coalesced stores instead of one thread looping over it.  Small rects (<= 4 tiles) stay per-thread.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
duplicate_kernel(int P, int gx, int gy, const int32_t* __restrict__ radii, const GeomRec* __restrict__ rec,
                 const float* __restrict__ depths, const uint32_t* __restrict__ offsets, uint64_t* __restrict__ keys,
                 uint32_t* __restrict__ vals, uint32_t capacity) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  uint32_t x0 = 0, y0 = 0, x1 = 0, y1 = 0, off = 0, dbits = 0;
  int radius = 0;
  if (idx < P) radius = radii[idx];
  if (radius > 0) {
    const float2 xy = *reinterpret_cast<const float2*>(rec + idx);
    get_rect(xy.x, xy.y, radius, gx, gy, x0, y0, x1, y1);
    off = idx == 0 ? 0u : offsets[idx - 1];
    dbits = __float_as_uint(depths[idx]);
  }
  const uint32_t w = x1 - x0;
  const uint32_t n = w * (y1 - y0);
  constexpr uint32_t SMALL = 4;
  if (n > 0 && n <= SMALL) {
    for (uint32_t i = 0; i < n; i++) {
      const uint32_t y = y0 + i / w, x = x0 + i % w;
      const uint32_t o = off + i;
      if (o < capacity) {
        keys[o] = ((uint64_t)(y * gx + x) << 32) | dbits;
        vals[o] = (uint32_t)idx;
      }
    }
  }
  uint32_t big = __ballot_sync(0xffffffffu, n > SMALL);
  while (big) {
    const int src = __ffs(big) - 1;
    big &= big - 1;
    const uint32_t bx0 = __shfl_sync(0xffffffffu, x0, src), by0 = __shfl_sync(0xffffffffu, y0, src);
    const uint32_t bw = __shfl_sync(0xffffffffu, w, src), bn = __shfl_sync(0xffffffffu, n, src);
    const uint32_t boff = __shfl_sync(0xffffffffu, off, src), bd = __shfl_sync(0xffffffffu, dbits, src);
    const uint32_t bidx = __shfl_sync(0xffffffffu, (uint32_t)idx, src);
    for (uint32_t i = lane; i < bn; i += 32) {
      const uint32_t y = by0 + i / bw, x = bx0 + i % bw;
      const uint32_t o = boff + i;
      if (o < capacity) {
        keys[o] = ((uint64_t)(y * gx + x) << 32) | bd;
        vals[o] = bidx;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------------------
// K5': identifyTileRanges (rasterizer_impl.cu:116-138).  ranges must be zero on entry (cudaMemset, :313).
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
tile_ranges_kernel(uint32_t L, const uint64_t* __restrict__ keys, uint2* __restrict__ ranges) {
  const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= L) return;
  const uint32_t cur = (uint32_t)(keys[idx] >> 32);
  if (idx == 0) {
    ranges[cur].x = 0;
  } else {
    const uint32_t prev = (uint32_t)(keys[idx - 1] >> 32);
    if (cur != prev) {
      ranges[prev].y = idx;
      ranges[cur].x = idx;
    }
  }
  if (idx == L - 1) ranges[cur].y = L;
}

}  // namespace

void launch_scan_tiles(int P, const uint32_t* tiles, uint32_t* offsets, uint64_t* status, uint32_t* ctrl,
                       cudaStream_t s) {
  if (P <= 0) return;
  const int blocks = (P + SCAN_TILE - 1) / SCAN_TILE;
  scan_tiles_kernel<<<blocks, SCAN_THREADS, 0, s>>>(P, tiles, offsets, status, ctrl);
  count_launch();
}

void launch_duplicate(int P, int grid_x, int grid_y, const int32_t* radii, GeomView g, uint64_t* keys, uint32_t* vals,
                      int64_t capacity, cudaStream_t s) {
  if (P <= 0) return;
  duplicate_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, grid_x, grid_y, radii, g.rec, g.depth, g.offsets, keys, vals,
                                                   (uint32_t)capacity);
  count_launch();
}

void launch_tile_ranges(int64_t R, const uint64_t* keys, uint2* ranges, cudaStream_t s) {
  if (R <= 0) return;
  tile_ranges_kernel<<<(unsigned)((R + 255) / 256), 256, 0, s>>>((uint32_t)R, keys, ranges);
  count_launch();
}

}  // namespace lsr
