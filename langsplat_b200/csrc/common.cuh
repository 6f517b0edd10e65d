// common.cuh — shared definitions for the sm_100a rasterizer kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

#include "../../include/lsr_api.h"

namespace lsr {

constexpr int TILE_X = 16, TILE_Y = 16;  // reference: BLOCK_X/BLOCK_Y, RAST/cuda_rasterizer/config.h:17-18
constexpr int TILE_PIX = TILE_X * TILE_Y;

// render_bwd accumulates pair-weighted moments instead of the six geometric gradient sums, preprocess_bwd applies the
// per-Gaussian coefficients (see render_bwd.cu).  0 restores the direct sums (A/B measurements).
#ifndef LSR_TILE_ORDER
#define LSR_TILE_ORDER 1  // blend kernels take the tiles heaviest first (shorter tail of the last wave)
#endif
#ifndef LSR_FWD_WEIGHT
#define LSR_FWD_WEIGHT 0  // forward blend: one weight alpha * T per pair instead of the reference's per-channel order (A/B)
#endif
#ifndef LSR_FWD_PK
#define LSR_FWD_PK 1  // forward blend: both pixels of a thread through packed fp32 (FFMA2) instructions
#endif
#ifndef LSR_MOMENTS
#define LSR_MOMENTS 1
#endif
#ifndef LSR_BWD_IEEE_DIV
#define LSR_BWD_IEEE_DIV 0  // 1: walk T back with IEEE divisions like the reference (A/B of the gradient error)
#endif

// One 48-byte record per Gaussian holds everything the blend kernels gather per instance, so a gather touches
// two 32-B sectors instead of the reference's four to six (means2D, conic_opacity, rgb, feature are separate
// arrays there: rasterizer_impl.cu:155-170).
struct __align__(16) GeomRec {
  float x, y, cx, cy;   // pixel centre, conic.x, conic.y
  float cz, o, r, g;    // conic.z, opacity, colour r g
  float b, f0, f1, f2;  // colour b, language feature
};
static_assert(sizeof(GeomRec) == 48, "GeomRec must be 48 bytes");

struct GeomView {
  GeomRec* rec;
  float* depth;
  uint32_t* tiles;
  uint32_t* offsets;
  float* cov3D;
  uint8_t* clamped;
};
struct ImageView {
  float* final_T;
  uint32_t* n_contrib;
  uint2* ranges;
  uint32_t* tile_order;  // tiles by descending list length: the blend kernels' CTA -> tile map (heaviest first)
  uint32_t* status;      // [LSR_STATUS_WORDS] what the forward found out on the device (lsr_api.h), kept per call
};

__host__ __device__ inline size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

// Function attributes (MaxDynamicSharedMemorySize) are per DEVICE: a process that drives several GPUs must set them
// on each one.  run(f) calls f() the first time it is reached with a given current device (idempotent work only:
// two threads racing here both run f, which is harmless).
struct DeviceOnce {
  std::atomic<uint64_t> seen{0};
  template <class F>
  void run(F&& f) {
    int dev = 0;
    cudaGetDevice(&dev);
    const uint64_t bit = 1ull << (dev & 63);
    if (seen.load(std::memory_order_acquire) & bit) return;
    f();
    seen.fetch_or(bit, std::memory_order_release);
  }
};

inline GeomView geom_view(void* base, int P) {
  char* p = (char*)base;
  GeomView g;
  size_t n = (size_t)(P > 0 ? P : 0);
  g.rec = (GeomRec*)p;       p += align_up(n * sizeof(GeomRec));
  g.depth = (float*)p;       p += align_up(n * 4);
  g.tiles = (uint32_t*)p;    p += align_up(n * 4);
  g.offsets = (uint32_t*)p;  p += align_up(n * 4);
  g.cov3D = (float*)p;       p += align_up(n * 24);
  g.clamped = (uint8_t*)p;   p += align_up(n);
  return g;
}
inline size_t geom_bytes(int P) {
  size_t n = (size_t)(P > 0 ? P : 0);
  return align_up(n * sizeof(GeomRec)) + 3 * align_up(n * 4) + align_up(n * 24) + align_up(n) + 256;
}
inline ImageView image_view(void* base, int W, int H) {
  char* p = (char*)base;
  size_t N = (size_t)W * H;
  size_t T = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
  ImageView v;
  v.final_T = (float*)p;       p += align_up(N * 4);
  v.n_contrib = (uint32_t*)p;  p += align_up(N * 4);
  v.ranges = (uint2*)p;        p += align_up(T * 8);
  v.tile_order = (uint32_t*)p; p += align_up(T * 4);
  v.status = (uint32_t*)p;     p += 256;
  return v;
}
inline size_t image_bytes(int W, int H) {
  size_t N = (size_t)W * H;
  size_t T = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
  return 2 * align_up(N * 4) + align_up(T * 8) + align_up(T * 4) + 256 + 256;
}

// ---- radix sort geometry (radix_sort.cu) ----
constexpr int SORT_RADIX_BITS = 8;
constexpr int SORT_BINS = 1 << SORT_RADIX_BITS;
constexpr int SORT_THREADS = 256;
#ifndef LSR_SORT_ITEMS
#define LSR_SORT_ITEMS 16  // GPU call r2y, C2 sort stage: 12 -> 136 us, 16 -> 132, 20 -> 159, 24 -> 174 (occupancy falls)
#endif
constexpr int SORT_ITEMS = LSR_SORT_ITEMS;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;  // keys per CTA per pass
inline int sort_items_for(int64_t) { return SORT_ITEMS; }  // keys per thread (the kernel is templated on it)
constexpr int SORT_MAX_PASSES = 8;

struct SortWork {  // carved from the `work` buffer
  uint32_t* hist;      // [SORT_MAX_PASSES][SORT_BINS] global digit histograms (raw counts)
  uint32_t* tickets;   // [SORT_MAX_PASSES] dynamic tile ids
  uint32_t* lookback;  // [SORT_MAX_PASSES][nblocks][SORT_BINS]
  size_t zero_bytes;   // bytes from hist to the end that must be zero before a sort
};
inline size_t sort_nblocks(int64_t n) { return (size_t)((n + SORT_TILE - 1) / SORT_TILE) + 1; }
inline size_t sort_work_bytes(int64_t n) {
  return align_up(SORT_MAX_PASSES * SORT_BINS * 4) + align_up(SORT_MAX_PASSES * 4) +
         align_up((size_t)SORT_MAX_PASSES * sort_nblocks(n) * SORT_BINS * 4) + 256;
}
inline SortWork sort_work_view(void* base, int64_t n) {
  char* p = (char*)base;
  SortWork w;
  w.hist = (uint32_t*)p;      p += align_up(SORT_MAX_PASSES * SORT_BINS * 4);
  w.tickets = (uint32_t*)p;   p += align_up(SORT_MAX_PASSES * 4);
  w.lookback = (uint32_t*)p;  p += align_up((size_t)SORT_MAX_PASSES * sort_nblocks(n) * SORT_BINS * 4);
  w.zero_bytes = (size_t)(p - (char*)base);
  return w;
}

// ---- scan workspace (binning.cu) ----
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;
inline size_t scan_nblocks(int P) { return (size_t)((P + SCAN_TILE - 1) / SCAN_TILE) + 1; }

#ifndef LSR_DUP_BLOCK
#define LSR_DUP_BLOCK 256  // 128 (twice the resident blocks) measured slower: 107 against 87 us at C2, 652 against 501 at C5 (r2y)
#endif
constexpr int DUP_BLOCK = LSR_DUP_BLOCK;  // Gaussians per block of the duplication kernel (binning.cu)
inline size_t dup_nblocks(int P) { return (size_t)((P + DUP_BLOCK - 1) / DUP_BLOCK) + 1; }

struct ScratchView {
  uint32_t* ctrl;  // [64] control words: [0] num_rendered, [1] scan ticket, [2] error flags
  uint64_t* scan_status;  // [max(scan_nblocks, dup_nblocks)] look-back words of the offset scan
  uint64_t* dwords_a;     // [P] depth-order sort of the Gaussians: packed (depth | index) words, ping
  uint64_t* dwords_b;     // [P] pong (sorted result)
  uint32_t* perm;         // [P] Gaussian indices in depth order
  void* sort_work_p;      // sort workspace for P elements
  uint64_t* keys_a;       // [cap] instance words (tile | index), ping-pong
  uint64_t* keys_b;       // [cap] sorted instance words end up here
  void* sort_work;        // sort workspace for cap elements
  size_t head_zero_bytes;  // ctrl + scan_status must be zero at the start of every forward
};
inline ScratchView scratch_view(void* base, int P, int64_t cap) {
  char* p = (char*)base;
  ScratchView s;
  size_t c = (size_t)(cap > 0 ? cap : 0);
  size_t n = (size_t)(P > 0 ? P : 0);
  s.ctrl = (uint32_t*)p;          p += 256;
  s.scan_status = (uint64_t*)p;   p += align_up(dup_nblocks(P) * 8);
  s.head_zero_bytes = (size_t)(p - (char*)base);
  s.dwords_a = (uint64_t*)p;      p += align_up(n * 8);
  s.dwords_b = (uint64_t*)p;      p += align_up(n * 8);
  s.perm = (uint32_t*)p;          p += align_up(n * 4);
  s.sort_work_p = (void*)p;       p += sort_work_bytes((int64_t)n);
  s.keys_a = (uint64_t*)p;        p += align_up(c * 8);
  s.keys_b = (uint64_t*)p;        p += align_up(c * 8);
  s.sort_work = (void*)p;
  return s;
}
inline size_t scratch_bytes(int P, int64_t cap) {
  size_t c = (size_t)(cap > 0 ? cap : 0);
  size_t n = (size_t)(P > 0 ? P : 0);
  return 256 + align_up(dup_nblocks(P) * 8) + 2 * align_up(n * 8) + align_up(n * 4) + sort_work_bytes((int64_t)n) +
         2 * align_up(c * 8) + sort_work_bytes(cap) + 256;
}

// ctrl words (zeroed at the start of every forward).  DEPTH_NEGMIN holds max(~bits) so that zero-init works.
// CTRL_DEPTH_BITS: width of the depth word's span, published by depth_words_kernel for the depth-sort passes.
enum CtrlWord { CTRL_NUM_RENDERED = 0, CTRL_SCAN_TICKET = 1, CTRL_ERROR = 2, CTRL_DEPTH_NEGMIN = 3, CTRL_DEPTH_MAX = 4,
                CTRL_DEPTH_BITS = 5, CTRL_DUP_TICKET = 6 };
constexpr uint32_t ERR_PREFILTERED = 1u, ERR_OVERFLOW = 2u;  // bits of ctrl[CTRL_ERROR] (= LSR_STATUS_FLAG_*)
constexpr int DEPTH_MAX_PASSES = 4;  // a 32-bit depth word in 8-bit digits

// Which bits the sort has to look at: digits are taken from (hi << lo_bits) | (lo - lo_min)  (radix_sort.cu).
// packed: the Gaussian id (id_bits) rides in the low bits of the 64-bit word, above it the compacted key; the sort
// then moves 8-byte words only.  Falls back to (key, value) pairs when hi_bits + lo_bits + id_bits > 64.
struct SortPlan {
  uint32_t lo_min;
  int lo_bits, hi_bits, total_bits, passes, id_bits;
  bool packed;
};
SortPlan make_sort_plan(uint32_t lo_min, uint32_t lo_max, int hi_bits, int id_bits);
SortPlan make_sort_plan_bits(int end_bit);
// allocation-free pair sort on the library's static device scratch (radix_sort.cu), n <= LSR_SORT_STATIC_MAX_N
cudaError_t sort_pairs_static(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                              int64_t n, int end_bit, cudaStream_t s);
// packed plan whose compacted key is `key_bits` wide (digits from bits [id_bits, id_bits + key_bits) of the word)
SortPlan make_sort_plan_packed(int key_bits, int id_bits);

// ---- launchers implemented in the .cu files ----
void launch_preprocess_fwd(const lsr_frame& f, const lsr_inputs& in, float focal_x, float focal_y, int32_t* radii,
                           GeomView g, uint32_t* ctrl, cudaStream_t s);
// packed (depth - lo_min | index) words of the P Gaussians, input of the depth-order sort, and that sort's digit
// histograms (ghist must be zero: sort_prepare)
// (the range of the depth word is read from ctrl on the device; its width is published in ctrl[CTRL_DEPTH_BITS])
void launch_depth_words(int P, const float* depths, const uint32_t* tiles, uint32_t* ctrl, int id_bits, uint64_t* words,
                        uint32_t* ghist, cudaStream_t s);
// duplication fused with the scan of the tile counts (look-back words in `status`, zero on entry; ticket in ctrl)
void launch_dup_scan(int P, int grid_x, int grid_y, const uint32_t* perm, const int32_t* radii, GeomView g,
                     uint64_t* words, int64_t capacity, const SortPlan& plan, uint64_t* status, uint32_t* ctrl,
                     uint32_t* ghist, cudaStream_t s);
// decodes the ranges marked by the last sort pass (encoded), orders the tiles, stores the frame's status words
void launch_finalize_tiles(int T, uint2* ranges, uint32_t* order, bool encoded, const uint32_t* ctrl,
                           int64_t capacity, uint32_t stamp, uint32_t* status, cudaStream_t s);
// NaN-fills a[0..na) and b[0..nb) if the status words report an instance-list overflow
void launch_poison_outputs(const uint32_t* status, float* a, size_t na, float* b, size_t nb, cudaStream_t s);
// inclusive scan of tiles[perm[i]] (perm == nullptr: of tiles[i], which also publishes the total in ctrl)
void launch_scan_tiles(int P, const uint32_t* tiles, const uint32_t* perm, uint32_t* offsets, uint64_t* status,
                       uint32_t* ctrl, cudaStream_t s);
void launch_duplicate(int P, int grid_x, int grid_y, const uint32_t* perm, const int32_t* radii, GeomView g,
                      uint64_t* words, int64_t capacity, const SortPlan& plan, uint32_t* ghist, cudaStream_t s);
// tile id of a sorted word = word >> tile_shift (32 for reference-format keys)
void launch_tile_ranges(int64_t R, const uint64_t* keys, int tile_shift, uint2* ranges, cudaStream_t s);
// order[] = the T tiles by descending list length (counting sort, one CTA); any R, also 0
void launch_tile_order(int T, uint2* ranges, uint32_t* order, cudaStream_t s);
// reference-format keys (tile << 32 | depth bits) from the sorted instance words (tests / debugging)
void launch_unpack_keys(int64_t R, const uint64_t* words, const float* depths, uint64_t* keys, int id_bits,
                        cudaStream_t s);
// hist_ready: the per-pass digit histograms are already in the work buffer (sort_prepare + duplication kernel)
void sort_prepare(void* work, int64_t n, const SortPlan& pl, cudaStream_t s);
// What only the device knows when the forward does not read anything back (all optional):
//  n_dev     element count (the `n` argument is then the bound the launch geometry and the workspace are sized for);
//  bits_dev  width of the packed key: max_passes passes are launched, those above the width return at once, the last
//            one below it writes the ids (the sorted words end up in keys_out OR keys_tmp: consume the ids only);
//  ranges    instance sort: the last pass marks the tile ranges (0xFF-filled on entry, decoded by finalize_tiles),
//            tile id = word >> tile_shift.
struct SortDyn {
  const uint32_t* n_dev = nullptr;
  const uint32_t* bits_dev = nullptr;
  int max_passes = 0;
  uint2* ranges = nullptr;
  int tile_shift = 0;
};
int sort_pairs(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
               uint64_t* keys_tmp, uint32_t* vals_tmp, int64_t n, const SortPlan& pl, bool hist_ready, void* work,
               cudaStream_t s, const SortDyn* dyn = nullptr);
void launch_render_fwd(int W, int H, const uint2* ranges, const uint32_t* order, const uint32_t* list, const GeomRec* rec, const float* bg,
                       bool include_feature, float* final_T, uint32_t* n_contrib, float* out_color,
                       float* out_feature, cudaStream_t s);
// feature-only re-blend of an already binned frame with another [P,3] language feature (records updated in place)
void launch_refeature_render(int P, int W, int H, const int32_t* radii, const float* lang, GeomRec* rec,
                             const uint2* ranges, const uint32_t* order, const uint32_t* list, float* out_feature,
                             cudaStream_t s);
void launch_render_bwd(int W, int H, const uint2* ranges, const uint32_t* order, const uint32_t* list, const GeomRec* rec, const float* bg,
                       bool include_feature, const float* final_T, const uint32_t* n_contrib, const float* dL_dcolor,
                       const float* dL_dfeature, uint32_t need_mask, float* dL_dmean2D, float* dL_dconic,
                       float* dL_dopacity, float* dL_dcolors, float* dL_dfeat, float* acc, cudaStream_t s);
// Packed accumulation record of the blend backward (lsr_grads::blend_acc): ACC_FLOATS floats per Gaussian, 64-byte
// aligned — (Mx, My, M0, 0) (Mxx, Mxy, Myy, 0) (dL_dcolour, 0) (dL_dfeature, 0); without LSR_NEED_GEOMETRY the first
// quarter holds the finished (dL_dmean2D.x, .y, 0, 0) and the second stays zero.  With `acc` given the kernel adds
// there (vector REDs) and leaves the five per-tensor buffers alone.
constexpr int ACC_FLOATS = 16;
bool blend_acc_supported();  // false when the one-phase kernel was selected (LSR_BWD_V1=1) or LSR_MOMENTS=0
// acc -> dL_dmean2D [P,3], dL_dcolors [P,3], dL_dfeat [P,3] (each optional) when no per-Gaussian kernel follows
void launch_unpack_acc(int P, const float* acc, float* dL_dmean2D, float* dL_dcolors, float* dL_dfeat, cudaStream_t s);
// the two-phase variant of the same kernel (render_bwd2.cu), same contract
void launch_render_bwd2(int W, int H, const uint2* ranges, const uint32_t* order, const uint32_t* list, const GeomRec* rec, const float* bg,
                        bool include_feature, const float* final_T, const uint32_t* n_contrib, const float* dL_dcolor,
                        const float* dL_dfeature, uint32_t need_mask, float* dL_dmean2D, float* dL_dconic,
                        float* dL_dopacity, float* dL_dcolors, float* dL_dfeat, float* acc, cudaStream_t s);
void launch_preprocess_bwd(const lsr_frame& f, const lsr_inputs& in, float focal_x, float focal_y,
                           const int32_t* radii, GeomView g, uint32_t need_mask, const lsr_grads& gr, const float* acc,
                           cudaStream_t s);
// dL_dsh [P,M,3] = sum over V views of basis(normalize(mean - campos_v)) (x) g_v, g: [V,P,3] clamp-masked dL_dcolor
void launch_sh_grad_views(int P, int D, int M, int V, const float* means3D, const float* campos, const float* g,
                          float* dL_dsh, cudaStream_t s);
// same, each view's factor / campos given by its own pointer (peer memory allowed); any V (chunks of LSR_MAX_VIEWS)
void launch_sh_grad_view_ptrs(int P, int D, int M, int V, const float* means3D, const float* const* g_views,
                              const float* const* campos_views, float* dL_dsh, cudaStream_t s);
// cross-GPU barrier on the stream (peer.cu): signals `epoch` into every peer's header, waits for all of them
// mode: 1 = signal, 2 = wait, 3 = both
void launch_peer_barrier(int rank, int world, void* const* headers, uint32_t epoch, int mode, cudaStream_t s);
void peer_set_timeout_ms(uint64_t ms);
uint64_t peer_timeout_ms();
// data[0..n) = NaN if a barrier of the owner of `own_buffer` ever timed out (sticky flag in its header)
void launch_peer_poison(const void* own_buffer, float* data, int64_t n, cudaStream_t s);
// out[r][i] = sum over ranks of in[.][i] for this rank's chunk of the n_floats-long vectors, for every rank r
void launch_peer_allreduce(int rank, int world, const float* const* in, float* const* out, int64_t n_floats,
                           cudaStream_t s);
void launch_activate_forward(int P, const float* scaling, const float* rotation, const float* opacity,
                             const float* feature, float* scales_out, float* rotations_out, float* opacities_out,
                             float* feature_out, cudaStream_t s);
void launch_activate_backward(int P, const float* scaling, const float* rotation, const float* opacity,
                              const float* feature, const float* g_scales, const float* g_rotations,
                              const float* g_opacities, const float* g_feature, float* d_scaling, float* d_rotation,
                              float* d_opacity, float* d_feature, cudaStream_t s);
void launch_feature_map(const void* seg, int seg_is_float, int64_t seg_row_stride, int W, int H, const float* table,
                        int K, int C, float* out, uint8_t* mask_u8, float* mask_f, int* bad, cudaStream_t s);
void decoder_set_trace(void* dev_buf);
size_t decoder_packed_bytes(const lsr_decoder* d, int precision);
void launch_decoder_pack(const lsr_decoder* d, int precision, void* packed, cudaStream_t s);
const char* decoder_tc_unsupported(const lsr_decoder* d, int precision);
cudaError_t launch_decode(const lsr_decoder* d, const float* in, int64_t n, int precision, float* out,
                          const float* phrases, int n_pos, int n_neg, float* relev, void* work, cudaStream_t s);
void launch_adam(int n, const lsr_adam_tensor* ts, int64_t step, double beta1, double beta2, double eps,
                 const uint8_t* visible, cudaStream_t s);
void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present, cudaStream_t s);

// loss.cu: fused (masked) L1 loss forward / backward on [C,N] images
cudaError_t launch_l1_forward(const float* pred, const float* gt, const float* mask, int C, int64_t N, float* loss,
                              cudaStream_t s);
cudaError_t launch_l1_backward(const float* pred, const float* gt, const float* mask, int C, int64_t N,
                               const float* grad_out, float* dL_dpred, cudaStream_t s);
// ... with the ground truth as the stored 8-bit image (gt = u / 255 formed in the kernel)
cudaError_t launch_l1_forward_u8(const float* pred, const uint8_t* gt, const float* mask, int C, int64_t N, float* loss,
                                 cudaStream_t s);
cudaError_t launch_l1_backward_u8(const float* pred, const uint8_t* gt, const float* mask, int C, int64_t N,
                                  const float* grad_out, float* dL_dpred, cudaStream_t s);

void count_launch(int n = 1);

}  // namespace lsr
