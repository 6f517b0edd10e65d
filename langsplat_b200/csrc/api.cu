// api.cu — the C ABI of liblsr.so (include/lsr_api.h): argument checks, buffer carving, stage orchestration.
// Orchestration replaces CudaRasterizer::Rasterizer::forward / backward / markVisible
// (RAST/cuda_rasterizer/rasterizer_impl.cu:141-153, 198-355, 359-463).
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <utility>
#include <vector>

#include "common.cuh"

namespace lsr {

static std::atomic<uint64_t> g_launches{0};
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

static thread_local char g_err[512] = "";
static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
#define LSR_CUDA(expr)                                                                                   \
  do {                                                                                                   \
    cudaError_t _e = (expr);                                                                             \
    if (_e != cudaSuccess) return fail(LSR_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(_e));    \
  } while (0)

static int stage_check(const lsr_frame* f, cudaStream_t s, const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess && f->debug) e = cudaStreamSynchronize(s);  // CHECK_CUDA(..., debug), auxiliary.h:166-173
  if (e != cudaSuccess) return fail(LSR_ERR_CUDA, "[CUDA ERROR] in stage %s: %s", what, cudaGetErrorString(e));
  return LSR_OK;
}
#define LSR_STAGE(what)                            \
  do {                                             \
    int _r = stage_check(frame, s, what);          \
    if (_r != LSR_OK) return _r;                   \
  } while (0)

// getHigherMsb, RAST/cuda_rasterizer/rasterizer_impl.cu:35-50
static uint32_t higher_msb(uint32_t n) {
  uint32_t msb = sizeof(n) * 4, step = msb;
  while (step > 1) {
    step /= 2;
    if (n >> msb) msb += step; else msb -= step;
  }
  if (n >> msb) msb++;
  return msb;
}

// ---- per-stage profiling with CUDA events on the caller's stream ----
enum { ST_PRE = 0, ST_SYNC, ST_DEPTH, ST_SCAN, ST_DUP, ST_SORT, ST_RANGES, ST_RENDER, ST_BZERO, ST_BRENDER, ST_BPRE,
       ST_COUNT };
static const char* kStageNames[ST_COUNT] = {"preprocess_fwd", "host_sync_gap", "depth_order", "scan", "duplicate",
                                            "sort", "tile_ranges", "render_fwd", "bwd_zero_fill", "render_bwd",
                                            "preprocess_bwd"};
static bool g_profile = false;
constexpr int N_EV_F = ST_RENDER - ST_PRE + 2;  // one event before each forward stage and one after the last
static cudaEvent_t g_ev_f[N_EV_F], g_ev_b[4];
static bool g_ev_init = false, g_have_f = false, g_have_b = false;
static void prof_init() {
  if (g_ev_init) return;
  for (auto& e : g_ev_f) cudaEventCreate(&e);
  for (auto& e : g_ev_b) cudaEventCreate(&e);
  g_ev_init = true;
}
static inline void prof(cudaEvent_t* arr, int i, cudaStream_t s) {
  if (g_profile) cudaEventRecord(arr[i], s);
}

// index width of the most recent forward on this thread (lsr_debug_sorted_keys decodes the instance words with it)
static thread_local int g_last_id_bits = 0;

static uint32_t* pinned_ctrl() {
  static thread_local uint32_t* p = nullptr;
  if (!p) cudaMallocHost(&p, 64 * sizeof(uint32_t));
  return p;
}

}  // namespace lsr

using namespace lsr;

#pragma GCC visibility push(default)
extern "C" {

size_t lsr_geom_bytes(int32_t P) { return geom_bytes(P); }
size_t lsr_image_bytes(int32_t width, int32_t height) { return image_bytes(width, height); }
size_t lsr_list_bytes(int64_t capacity) { return align_up((size_t)(capacity > 0 ? capacity : 0) * 4) + 256; }
size_t lsr_scratch_bytes(int32_t P, int32_t width, int32_t height, int64_t capacity) {
  (void)width; (void)height;
  return scratch_bytes(P, capacity);
}
size_t lsr_sort_work_bytes(int64_t n) { return sort_work_bytes(n); }

size_t lsr_array_offset(int which, int32_t P, int32_t width, int32_t height, int64_t capacity) {
  char* base = (char*)4096;  // never dereferenced
  GeomView g = geom_view(base, P);
  ImageView im = image_view(base, width, height);
  ScratchView sc = scratch_view(base, P, capacity);
  switch (which) {
    case LSR_ARR_GEOM_REC: return (char*)g.rec - base;
    case LSR_ARR_GEOM_DEPTH: return (char*)g.depth - base;
    case LSR_ARR_GEOM_TILES: return (char*)g.tiles - base;
    case LSR_ARR_GEOM_OFFSETS: return (char*)g.offsets - base;
    case LSR_ARR_GEOM_COV3D: return (char*)g.cov3D - base;
    case LSR_ARR_GEOM_CLAMPED: return (char*)g.clamped - base;
    case LSR_ARR_IMG_FINAL_T: return (char*)im.final_T - base;
    case LSR_ARR_IMG_NCONTRIB: return (char*)im.n_contrib - base;
    case LSR_ARR_IMG_RANGES: return (char*)im.ranges - base;
    case LSR_ARR_SCR_KEYS_UNSORTED: return (char*)sc.keys_a - base;
    case LSR_ARR_SCR_DEPTH_ORDER: return (char*)sc.perm - base;
    case LSR_ARR_SCR_KEYS_SORTED: return (char*)sc.keys_b - base;
    default: return (size_t)-1;
  }
}

const char* lsr_last_error(void) { return g_err; }
int lsr_version(void) { return LSR_VERSION; }
int lsr_blend_acc_supported(void) { return blend_acc_supported() ? 1 : 0; }
uint64_t lsr_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

void lsr_profile_enable(int on) {
  g_profile = on != 0;
  if (g_profile) prof_init();
  g_have_f = g_have_b = false;
}
const char* lsr_profile_stage_name(int i) { return (i >= 0 && i < ST_COUNT) ? kStageNames[i] : ""; }
int lsr_profile_read(float* ms, int n) {
  if (!g_profile) return 0;
  float v[ST_COUNT];
  for (int i = 0; i < ST_COUNT; i++) v[i] = 0.f;
  if (g_have_f) {
    cudaEventSynchronize(g_ev_f[N_EV_F - 1]);
    for (int i = 0; i < N_EV_F - 1; i++) cudaEventElapsedTime(&v[ST_PRE + i], g_ev_f[i], g_ev_f[i + 1]);
  }
  if (g_have_b) {
    cudaEventSynchronize(g_ev_b[3]);
    for (int i = 0; i < 3; i++) cudaEventElapsedTime(&v[ST_BZERO + i], g_ev_b[i], g_ev_b[i + 1]);
  }
  for (int i = 0; i < ST_COUNT && i < n; i++) ms[i] = v[i];
  return ST_COUNT;
}

int lsr_l1_loss_forward(const float* pred, const float* gt, const float* mask, int32_t C, int64_t N, float* loss,
                        void* stream) {
  if (C < 0 || N < 0) return fail(LSR_ERR_INVALID, "C < 0 or N < 0");
  if (!loss || ((int64_t)C * N > 0 && (!pred || !gt))) return fail(LSR_ERR_INVALID, "null pointer");
  LSR_CUDA(launch_l1_forward(pred, gt, mask, C, N, loss, (cudaStream_t)stream));
  return LSR_OK;
}

int lsr_l1_loss_backward(const float* pred, const float* gt, const float* mask, int32_t C, int64_t N,
                         const float* grad_out, float* dL_dpred, void* stream) {
  if (C < 0 || N < 0) return fail(LSR_ERR_INVALID, "C < 0 or N < 0");
  if ((int64_t)C * N > 0 && (!pred || !gt || !grad_out || !dL_dpred)) return fail(LSR_ERR_INVALID, "null pointer");
  LSR_CUDA(launch_l1_backward(pred, gt, mask, C, N, grad_out, dL_dpred, (cudaStream_t)stream));
  return LSR_OK;
}

int lsr_l1_loss_forward_u8(const float* pred, const uint8_t* gt, const float* mask, int32_t C, int64_t N, float* loss,
                           void* stream) {
  if (C < 0 || N < 0) return fail(LSR_ERR_INVALID, "C < 0 or N < 0");
  if (!loss || ((int64_t)C * N > 0 && (!pred || !gt))) return fail(LSR_ERR_INVALID, "null pointer");
  LSR_CUDA(launch_l1_forward_u8(pred, gt, mask, C, N, loss, (cudaStream_t)stream));
  return LSR_OK;
}

int lsr_l1_loss_backward_u8(const float* pred, const uint8_t* gt, const float* mask, int32_t C, int64_t N,
                            const float* grad_out, float* dL_dpred, void* stream) {
  if (C < 0 || N < 0) return fail(LSR_ERR_INVALID, "C < 0 or N < 0");
  if ((int64_t)C * N > 0 && (!pred || !gt || !grad_out || !dL_dpred)) return fail(LSR_ERR_INVALID, "null pointer");
  LSR_CUDA(launch_l1_backward_u8(pred, gt, mask, C, N, grad_out, dL_dpred, (cudaStream_t)stream));
  return LSR_OK;
}

int lsr_sh_gradient_from_views(int32_t P, int32_t M, int32_t D, int32_t n_views, const float* means3D,
                               const float* campos, const float* g, float* dL_dsh, void* stream) {
  if (P < 0 || M < 0 || n_views < 0 || D < 0 || D > 3) return fail(LSR_ERR_INVALID, "bad size / degree");
  if (M < (D + 1) * (D + 1)) return fail(LSR_ERR_INVALID, "sh_degree %d needs at least %d SH coefficients, got %d", D,
                                         (D + 1) * (D + 1), M);
  if (P > 0 && M > 0 && (!means3D || !dL_dsh || (n_views > 0 && (!campos || !g))))
    return fail(LSR_ERR_INVALID, "null pointer");
  launch_sh_grad_views(P, D, M, n_views, means3D, campos, g, dL_dsh, (cudaStream_t)stream);
  LSR_CUDA(cudaGetLastError());
  return LSR_OK;
}

int lsr_sh_gradient_from_view_ptrs(int32_t P, int32_t M, int32_t D, int32_t n_views, const float* means3D,
                                   const float* const* g_views, const float* const* campos_views, float* dL_dsh,
                                   void* stream) {
  if (P < 0 || M < 0 || n_views < 0 || D < 0 || D > 3) return fail(LSR_ERR_INVALID, "bad size / degree");
  if (M < (D + 1) * (D + 1)) return fail(LSR_ERR_INVALID, "sh_degree %d needs at least %d SH coefficients, got %d", D,
                                         (D + 1) * (D + 1), M);
  if (P > 0 && M > 0 && (!means3D || !dL_dsh || (n_views > 0 && (!g_views || !campos_views))))
    return fail(LSR_ERR_INVALID, "null pointer");
  for (int v = 0; v < n_views && P > 0 && M > 0; v++)
    if (!g_views[v] || !campos_views[v]) return fail(LSR_ERR_INVALID, "null pointer for view %d", v);
  launch_sh_grad_view_ptrs(P, D, M, n_views, means3D, g_views, campos_views, dL_dsh, (cudaStream_t)stream);
  LSR_CUDA(cudaGetLastError());
  return LSR_OK;
}

// ---- peer memory (CUDA IPC) ----
static_assert(sizeof(cudaIpcMemHandle_t) == LSR_PEER_HANDLE_BYTES, "IPC handle size");

int lsr_peer_alloc(size_t bytes, void** ptr, uint8_t* handle) {
  if (!ptr || !handle || bytes < LSR_PEER_HEADER_BYTES) return fail(LSR_ERR_INVALID, "bad argument");
  void* p = nullptr;
  LSR_CUDA(cudaMalloc(&p, bytes));
  cudaError_t e = cudaMemset(p, 0, bytes);
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    cudaFree(p);
    return fail(LSR_ERR_CUDA, "lsr_peer_alloc: %s", cudaGetErrorString(e));
  }
  memcpy(handle, &h, sizeof(h));
  *ptr = p;
  return LSR_OK;
}

int lsr_peer_open(const uint8_t* handle, void** ptr) {
  if (!handle || !ptr) return fail(LSR_ERR_INVALID, "null pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  LSR_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return LSR_OK;
}

int lsr_peer_close(void* ptr) {
  if (ptr) LSR_CUDA(cudaIpcCloseMemHandle(ptr));
  return LSR_OK;
}

int lsr_peer_free(void* ptr) {
  if (ptr) LSR_CUDA(cudaFree(ptr));
  return LSR_OK;
}

static int peer_sync(int32_t rank, int32_t world, void* const* headers, uint32_t epoch, int mode, void* stream) {
  if (world < 1 || world > LSR_MAX_PEERS || rank < 0 || rank >= world || !headers)
    return fail(LSR_ERR_INVALID, "bad rank / world (at most %d peers)", LSR_MAX_PEERS);
  for (int r = 0; r < world; r++)
    if (!headers[r]) return fail(LSR_ERR_INVALID, "null header for rank %d", r);
  launch_peer_barrier(rank, world, headers, epoch, mode, (cudaStream_t)stream);
  LSR_CUDA(cudaGetLastError());
  return LSR_OK;
}

int lsr_peer_barrier(int32_t rank, int32_t world, void* const* headers, uint32_t epoch, void* stream) {
  return peer_sync(rank, world, headers, epoch, 3, stream);
}
int lsr_peer_signal(int32_t rank, int32_t world, void* const* headers, uint32_t epoch, void* stream) {
  return peer_sync(rank, world, headers, epoch, 1, stream);
}
int lsr_peer_wait(int32_t rank, int32_t world, void* const* headers, uint32_t epoch, void* stream) {
  return peer_sync(rank, world, headers, epoch, 2, stream);
}

int lsr_peer_sync_auto(int32_t rank, int32_t world, void* const* headers, int32_t mode, void* stream) {
  if (mode < 1 || mode > 3) return fail(LSR_ERR_INVALID, "mode must be 1 (signal), 2 (wait) or 3 (barrier)");
  return peer_sync(rank, world, headers, 0u, mode | 4, stream);
}

int lsr_peer_allreduce(int32_t rank, int32_t world, const float* const* in, float* const* out, int64_t n_floats,
                       void* stream) {
  if (world < 1 || world > LSR_MAX_PEERS || rank < 0 || rank >= world || !in || !out)
    return fail(LSR_ERR_INVALID, "bad rank / world (at most %d peers)", LSR_MAX_PEERS);
  if (n_floats < 0 || (n_floats & 3)) return fail(LSR_ERR_INVALID, "n_floats must be a non-negative multiple of 4");
  for (int r = 0; r < world; r++)
    if (!in[r] || !out[r] || ((((uintptr_t)in[r]) | ((uintptr_t)out[r])) & 15))
      return fail(LSR_ERR_INVALID, "buffer of rank %d is null or not 16-byte aligned", r);
  launch_peer_allreduce(rank, world, in, out, n_floats, (cudaStream_t)stream);
  LSR_CUDA(cudaGetLastError());
  return LSR_OK;
}

int lsr_peer_set_timeout_ms(uint64_t ms) {
  if (ms == 0) return fail(LSR_ERR_INVALID, "time-out must be at least 1 ms");
  peer_set_timeout_ms(ms);
  return LSR_OK;
}

int lsr_peer_poison_on_timeout(const void* own_buffer, float* data, int64_t n_floats, void* stream) {
  if (!own_buffer || (n_floats > 0 && !data) || n_floats < 0) return fail(LSR_ERR_INVALID, "bad argument");
  launch_peer_poison(own_buffer, data, n_floats, (cudaStream_t)stream);
  LSR_CUDA(cudaGetLastError());
  return LSR_OK;
}

int lsr_peer_status(const void* own_buffer, void* stream) {
  if (!own_buffer) return fail(LSR_ERR_INVALID, "null pointer");
  uint32_t flag = 0;
  LSR_CUDA(cudaMemcpyAsync(&flag, static_cast<const uint8_t*>(own_buffer) + LSR_MAX_PEERS * sizeof(uint32_t),
                           sizeof(flag), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  LSR_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  if (flag) return fail(LSR_ERR_CUDA, "peer barrier timed out (a rank did not arrive within %llu ms)",
                        (unsigned long long)peer_timeout_ms());
  return LSR_OK;
}

int lsr_activate_forward(int32_t P, const float* scaling, const float* rotation, const float* opacity,
                         const float* feature, float* scales_out, float* rotations_out, float* opacities_out,
                         float* feature_out, void* stream) {
  if (P < 0) return fail(LSR_ERR_INVALID, "P < 0");
  if (P == 0) return LSR_OK;
  if ((scaling && !scales_out) || (rotation && !rotations_out) || (opacity && !opacities_out) ||
      (feature && !feature_out))
    return fail(LSR_ERR_INVALID, "an input is given without its output");
  if (rotation && ((((uintptr_t)rotation) | ((uintptr_t)rotations_out)) & 15))
    return fail(LSR_ERR_INVALID, "rotation tensors must be 16-byte aligned");
  launch_activate_forward(P, scaling, rotation, opacity, feature, scales_out, rotations_out, opacities_out,
                          feature_out, (cudaStream_t)stream);
  LSR_CUDA(cudaGetLastError());
  return LSR_OK;
}

int lsr_activate_backward(int32_t P, const float* scaling, const float* rotation, const float* opacity,
                          const float* feature, const float* g_scales, const float* g_rotations,
                          const float* g_opacities, const float* g_feature, float* d_scaling, float* d_rotation,
                          float* d_opacity, float* d_feature, void* stream) {
  if (P < 0) return fail(LSR_ERR_INVALID, "P < 0");
  if (P == 0) return LSR_OK;
  if ((d_scaling && (!scaling || !g_scales)) || (d_rotation && (!rotation || !g_rotations)) ||
      (d_opacity && (!opacity || !g_opacities)) || (d_feature && (!feature || !g_feature)))
    return fail(LSR_ERR_INVALID, "a gradient is requested without its raw input or upstream gradient");
  if (d_rotation && ((((uintptr_t)rotation) | ((uintptr_t)g_rotations) | ((uintptr_t)d_rotation)) & 15))
    return fail(LSR_ERR_INVALID, "rotation tensors must be 16-byte aligned");
  launch_activate_backward(P, scaling, rotation, opacity, feature, g_scales, g_rotations, g_opacities, g_feature,
                           d_scaling, d_rotation, d_opacity, d_feature, (cudaStream_t)stream);
  LSR_CUDA(cudaGetLastError());
  return LSR_OK;
}

int lsr_language_feature_map(const void* seg, int32_t seg_is_float, int64_t seg_row_stride, int32_t width,
                             int32_t height, const float* table, int32_t K, int32_t C, float* out, uint8_t* mask_u8,
                             float* mask_f, int32_t* bad, void* stream) {
  if (width < 0 || height < 0 || K < 1 || C < 1 || seg_row_stride < width)
    return fail(LSR_ERR_INVALID, "bad size (K and C must be >= 1, seg_row_stride >= width)");
  if ((int64_t)width * height == 0) return LSR_OK;
  if (!seg || !table || !out) return fail(LSR_ERR_INVALID, "null pointer");
  launch_feature_map(seg, seg_is_float, seg_row_stride, width, height, table, K, C, out, mask_u8, mask_f, bad,
                     (cudaStream_t)stream);
  LSR_CUDA(cudaGetLastError());
  return LSR_OK;
}

int lsr_adam_step(int32_t n_tensors, const lsr_adam_tensor* tensors, int64_t step, double beta1, double beta2,
                  double eps, const uint8_t* visible, void* stream) {
  if (n_tensors < 0 || n_tensors > LSR_ADAM_MAX_TENSORS)
    return fail(LSR_ERR_INVALID, "at most %d tensors per call", LSR_ADAM_MAX_TENSORS);
  if (step < 1) return fail(LSR_ERR_INVALID, "step must be >= 1 (the count after this step)");
  if (!(beta1 >= 0.0 && beta1 < 1.0) || !(beta2 >= 0.0 && beta2 < 1.0) || !(eps >= 0.0))
    return fail(LSR_ERR_INVALID, "bad betas / eps");
  if (n_tensors > 0 && !tensors) return fail(LSR_ERR_INVALID, "null pointer");
  for (int i = 0; i < n_tensors; i++) {
    const lsr_adam_tensor& t = tensors[i];
    if (t.numel < 0) return fail(LSR_ERR_INVALID, "tensor %d: numel < 0", i);
    if (t.numel > 0 && (!t.param || !t.grad || !t.exp_avg || !t.exp_avg_sq))
      return fail(LSR_ERR_INVALID, "tensor %d: null pointer", i);
    if (visible && t.numel > 0 && (t.row_width < 1 || t.numel % t.row_width))
      return fail(LSR_ERR_INVALID, "tensor %d: numel is not a multiple of row_width", i);
  }
  launch_adam(n_tensors, tensors, step, beta1, beta2, eps, visible, (cudaStream_t)stream);
  LSR_CUDA(cudaGetLastError());
  return LSR_OK;
}

static int decoder_check(const lsr_decoder* d) {
  if (!d) return fail(LSR_ERR_INVALID, "null decoder");
  if (d->n_layers < 1 || d->n_layers > LSR_DECODER_MAX_LAYERS)
    return fail(LSR_ERR_INVALID, "decoder needs 1..%d layers", LSR_DECODER_MAX_LAYERS);
  for (int l = 0; l <= d->n_layers; l++)
    if (d->dims[l] < 1 || d->dims[l] > 512) return fail(LSR_ERR_INVALID, "decoder widths must be in 1..512");
  for (int l = 0; l < d->n_layers; l++)
    if (!d->weight[l] || !d->bias[l]) return fail(LSR_ERR_INVALID, "layer %d: null pointer", l);
  return LSR_OK;
}

int lsr_debug_decode_trace(void* device_buffer) {
  decoder_set_trace(device_buffer);
  return LSR_OK;
}

static bool decoder_tc_precision(int32_t precision) {
  return precision == LSR_DECODE_TF32 || precision == LSR_DECODE_F16 || precision == LSR_DECODE_BF16X2;
}

size_t lsr_decoder_packed_bytes(const lsr_decoder* dec, int32_t precision) {
  if (decoder_check(dec) != LSR_OK) return 0;
  if (!decoder_tc_precision(precision)) { fail(LSR_ERR_INVALID, "only the tensor-core variants use packed weights"); return 0; }
  return decoder_packed_bytes(dec, precision);
}

int lsr_decoder_pack(const lsr_decoder* dec, int32_t precision, void* packed, void* stream) {
  if (int rc = decoder_check(dec)) return rc;
  if (!decoder_tc_precision(precision)) return fail(LSR_ERR_INVALID, "only the tensor-core variants use packed weights");
  if (!packed || ((uintptr_t)packed & 15)) return fail(LSR_ERR_INVALID, "packed must be a 16-byte aligned device pointer");
  launch_decoder_pack(dec, precision, packed, (cudaStream_t)stream);
  LSR_CUDA(cudaGetLastError());
  return LSR_OK;
}

size_t lsr_decode_work_bytes(void) { return 32768; }

int lsr_decode(const lsr_decoder* dec, const float* in, int64_t n, int32_t precision, float* out, const float* phrases,
               int32_t n_pos, int32_t n_neg, float* relevancy, void* work, size_t work_bytes, void* stream) {
  if (int rc = decoder_check(dec)) return rc;
  if (n < 0) return fail(LSR_ERR_INVALID, "n < 0");
  if (precision != LSR_DECODE_FP32 && !decoder_tc_precision(precision)) return fail(LSR_ERR_INVALID, "unknown precision");
  if (relevancy) {
    if (!phrases) return fail(LSR_ERR_INVALID, "relevancy needs phrase embeddings");
    if (n_pos < 1 || n_neg < 1 || n_pos + n_neg > LSR_DECODER_MAX_PHRASES)
      return fail(LSR_ERR_INVALID, "need n_pos >= 1, n_neg >= 1, n_pos + n_neg <= %d", LSR_DECODER_MAX_PHRASES);
  } else {
    phrases = nullptr;
    n_pos = n_neg = 0;
  }
  if (!out && !relevancy) return fail(LSR_ERR_INVALID, "nothing to compute (out and relevancy are both null)");
  if (n == 0) return LSR_OK;
  if (!in) return fail(LSR_ERR_INVALID, "null pointer");
  if (decoder_tc_precision(precision)) {
    if (const char* why = decoder_tc_unsupported(dec, precision)) return fail(LSR_ERR_INVALID, "tensor-core decode: %s", why);
    if (out && ((uintptr_t)out & 15)) return fail(LSR_ERR_INVALID, "out must be 16-byte aligned");
    if (relevancy && (!work || ((uintptr_t)work & 15) || work_bytes < lsr_decode_work_bytes()))
      return fail(LSR_ERR_INVALID, "tensor-core relevancy needs a 16-byte aligned work buffer of lsr_decode_work_bytes()");
  }
  LSR_CUDA(launch_decode(dec, in, n, precision, out, phrases, n_pos, n_neg, relevancy, work, (cudaStream_t)stream));
  return LSR_OK;
}

int lsr_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream) {
  (void)projmatrix;  // the reference's in_frustum computes p_proj but only tests p_view.z (auxiliary.h:139-164)
  if (P < 0) return fail(LSR_ERR_INVALID, "P < 0");
  if (P == 0) return LSR_OK;
  if (!means3D || !viewmatrix || !present) return fail(LSR_ERR_INVALID, "null pointer");
  launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream);
  LSR_CUDA(cudaGetLastError());
  return LSR_OK;
}

int lsr_debug_sorted_keys(const void* scratch, const void* geom, int32_t P, int64_t capacity, int64_t num_rendered,
                          uint64_t* keys_out, void* stream) {
  if (num_rendered == 0) return LSR_OK;
  if (!scratch || !geom || !keys_out || num_rendered < 0 || num_rendered > capacity || g_last_id_bits <= 0)
    return fail(LSR_ERR_INVALID, "bad argument");
  ScratchView sc = scratch_view(const_cast<void*>(scratch), P, capacity);
  GeomView g = geom_view(const_cast<void*>(geom), P);
  launch_unpack_keys(num_rendered, sc.keys_b, g.depth, keys_out, g_last_id_bits, (cudaStream_t)stream);
  LSR_CUDA(cudaGetLastError());
  return LSR_OK;
}

int lsr_sort_pairs(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                   uint64_t* keys_tmp, uint32_t* vals_tmp, int64_t n, int end_bit, void* work, size_t work_bytes,
                   void* stream) {
  if (n < 0) return fail(LSR_ERR_INVALID, "n < 0");
  if (n == 0) return LSR_OK;
  if (n >= (1ll << 30)) return fail(LSR_ERR_INVALID, "n must be < 2^30");
  if (!keys_in || !vals_in || !keys_out || !vals_out || !keys_tmp || !vals_tmp || !work)
    return fail(LSR_ERR_INVALID, "null pointer");
  if (work_bytes < sort_work_bytes(n)) return fail(LSR_ERR_CAPACITY, "sort work buffer too small");
  sort_pairs(keys_in, vals_in, keys_out, vals_out, keys_tmp, vals_tmp, n, make_sort_plan_bits(end_bit), false, work,
             (cudaStream_t)stream);
  LSR_CUDA(cudaGetLastError());
  return LSR_OK;
}

int lsr_sort_pairs_static(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                          int64_t n, int end_bit, void* stream) {
  if (n < 0) return fail(LSR_ERR_INVALID, "n < 0");
  if (n == 0) return LSR_OK;
  if (n > (int64_t)LSR_SORT_STATIC_MAX_N)
    return fail(LSR_ERR_CAPACITY, "lsr_sort_pairs_static: n = %lld exceeds LSR_SORT_STATIC_MAX_N = %d pairs (the size "
                "of the library's static scratch); use lsr_sort_pairs with caller-provided buffers",
                (long long)n, (int)LSR_SORT_STATIC_MAX_N);
  if (!keys_in || !vals_in || !keys_out || !vals_out) return fail(LSR_ERR_INVALID, "null pointer");
  LSR_CUDA(sort_pairs_static(keys_in, vals_in, keys_out, vals_out, n, end_bit, (cudaStream_t)stream));
  return LSR_OK;
}

static int check_frame(const lsr_frame* f, const lsr_inputs* in) {
  if (!f || !in) return fail(LSR_ERR_INVALID, "null frame / inputs");
  if (f->P < 0 || f->width <= 0 || f->height <= 0) return fail(LSR_ERR_INVALID, "bad P / image size");
  if (f->P == 0) return LSR_OK;
  if (!in->means3D) return fail(LSR_ERR_INVALID, "means3D must have dimensions (num_points, 3)");
  if (!in->opacities || !f->bg || !f->viewmatrix || !f->projmatrix || !f->campos)
    return fail(LSR_ERR_INVALID, "null opacities / bg / viewmatrix / projmatrix / campos");
  if ((in->shs == nullptr) == (in->colors_precomp == nullptr))
    return fail(LSR_ERR_INVALID, "Please provide excatly one of either SHs or precomputed colors!");
  const bool sr = in->scales != nullptr && in->rotations != nullptr;
  if (sr == (in->cov3D_precomp != nullptr) || ((in->scales != nullptr) != (in->rotations != nullptr)))
    return fail(LSR_ERR_INVALID,
                "Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!");
  if (in->shs && (f->D < 0 || f->D > 3 || f->M < (f->D + 1) * (f->D + 1)))
    return fail(LSR_ERR_INVALID, "sh_degree %d needs at least %d SH coefficients, got %d", f->D,
                (f->D + 1) * (f->D + 1), f->M);
  if (f->include_feature && !in->language_feature)
    return fail(LSR_ERR_INVALID, "include_feature is set but language_feature_precomp is missing");
  // the kernels read quaternions as one 128-bit load per Gaussian: a contiguous view that starts at an odd element
  // of its storage (x[1:]) would fault, so say so here instead (the Python layer re-packs such tensors)
  if (in->rotations && (((uintptr_t)in->rotations) & 15))
    return fail(LSR_ERR_INVALID, "rotations must be 16-byte aligned");
  return LSR_OK;
}

// One device pipeline serves both entry points.  Nothing after the preprocess kernel needs the instance count R or
// the depth range on the host: the kernels read them from the control words (binning.cu, radix_sort.cu), launch
// geometry comes from a bound on R.  lsr_forward reads R back once after the preprocess kernel, like the reference
// (rasterizer_impl.cu:284), because its contract is to report R and to refuse a list that is too small BEFORE anything
// is blended; its bound is then R itself.  lsr_forward_nosync skips that read: the bound is the list capacity, an
// overflow is flagged on the device and the images are NaN-filled.
static int forward_impl(const lsr_frame* frame, const lsr_inputs* in, float* out_color, float* out_feature,
                        int32_t* radii, void* geom, void* image, void* list, int64_t list_capacity, void* scratch,
                        size_t scratch_nbytes, int64_t* num_rendered, bool sync, uint32_t stamp,
                        uint32_t* status_host, cudaStream_t s) {
  int rc = check_frame(frame, in);
  if (rc != LSR_OK) return rc;
  if (!out_color || !image || (sync && !num_rendered) || (frame->include_feature && !out_feature))
    return fail(LSR_ERR_INVALID, "null output pointer");
  const int P = frame->P, W = frame->width, H = frame->height;
  const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
  const int T = gx * gy;
  const size_t N = (size_t)W * H;
  ImageView img = image_view(image, W, H);
  if (num_rendered) *num_rendered = 0;
  if (g_profile) g_have_f = false;
  if (P == 0) {
    // rasterize_points.cu:91: the reference returns zero-filled images for an empty scene (blending empty ranges
    // would give colour = bg, which differs), so reproduce the zero fill
    LSR_CUDA(cudaMemsetAsync(img.ranges, 0, (size_t)T * sizeof(uint2), s));
    LSR_CUDA(cudaMemsetAsync(img.status, 0, LSR_STATUS_WORDS * sizeof(uint32_t), s));
    LSR_CUDA(cudaMemsetAsync(out_color, 0, 3 * N * sizeof(float), s));
    if (frame->include_feature) LSR_CUDA(cudaMemsetAsync(out_feature, 0, 3 * N * sizeof(float), s));
    if (g_profile) {
      for (int i = 0; i <= 8; i++) prof(g_ev_f, i, s);
      g_have_f = true;
    }
    if (status_host) {
      // (the stamp travels through a staging word of the image state so that the copy stays stream-ordered)
      LSR_CUDA(cudaMemcpyAsync(img.status + LSR_STATUS_STAMP, &stamp, sizeof(stamp), cudaMemcpyHostToDevice, s));
      LSR_CUDA(cudaMemcpyAsync(status_host, img.status, LSR_STATUS_WORDS * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    }
    return LSR_OK;
  }
  if (!radii || !geom || !scratch) return fail(LSR_ERR_INVALID, "null radii / geom / scratch");
  if (list_capacity < 0 || list_capacity >= (1ll << 30)) return fail(LSR_ERR_INVALID, "list_capacity out of range");
  if (list_capacity > 0 && !list) return fail(LSR_ERR_INVALID, "null list buffer");
  if (scratch_nbytes < scratch_bytes(P, list_capacity))
    return fail(LSR_ERR_INVALID, "scratch buffer smaller than lsr_scratch_bytes(P, capacity)");
  GeomView g = geom_view(geom, P);
  ScratchView sc = scratch_view(scratch, P, list_capacity);
  // focal lengths exactly as the reference computes them on the host in f32 (rasterizer_impl.cu:225-226)
  const float focal_y = H / (2.0f * frame->tanfovy);
  const float focal_x = W / (2.0f * frame->tanfovx);
  int id_bits = 1;
  while (id_bits < 32 && (((uint32_t)P - 1u) >> id_bits) != 0) id_bits++;
  // The reference sorts R keys on bits [0, 32 + getHigherMsb(T)) (rasterizer_impl.cu:303,311).  Same order, less
  // traffic (binning.cu): (1) stable sort of the P Gaussians by depth — the depth word only spans [min, max], so
  // bitlength(max - min) bits suffice, a width only the device knows; (2) instances emitted in that order;
  // (3) stable sort of the R instance words on the tile bits only.
  const SortPlan dplan = make_sort_plan_packed(32, id_bits);  // worst case: DEPTH_MAX_PASSES, the device skips the rest
  const SortPlan plan = make_sort_plan_packed((int)higher_msb((uint32_t)T), id_bits);
  g_last_id_bits = id_bits;

  LSR_CUDA(cudaMemsetAsync(scratch, 0, sc.head_zero_bytes, s));
  LSR_CUDA(cudaMemsetAsync(img.ranges, 0xFF, (size_t)T * sizeof(uint2), s));  // (start, ~end) under min-operations
  prof(g_ev_f, 0, s);
  launch_preprocess_fwd(*frame, *in, focal_x, focal_y, radii, g, sc.ctrl, s);
  LSR_STAGE("preprocess_fwd");
  prof(g_ev_f, 1, s);
  int64_t bound = list_capacity;  // what the launch geometry of the R-sized kernels is sized for
  if (sync) {
    uint32_t* h = pinned_ctrl();
    LSR_CUDA(cudaMemcpyAsync(h, sc.ctrl, 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    LSR_CUDA(cudaStreamSynchronize(s));
    const int64_t R = (int64_t)h[CTRL_NUM_RENDERED];
    *num_rendered = R;
    if (h[CTRL_ERROR] & ERR_PREFILTERED)
      return fail(LSR_ERR_PREFILTERED, "Point is filtered although prefiltered is set. This shouldn't happen!");
    if (R > list_capacity) return fail(LSR_ERR_CAPACITY, "num_rendered %lld exceeds list capacity %lld",
                                       (long long)R, (long long)list_capacity);
    bound = R;
  }
  prof(g_ev_f, 2, s);
  {
    sort_prepare(sc.sort_work_p, P, dplan, s);
    launch_depth_words(P, g.depth, g.tiles, sc.ctrl, id_bits, sc.dwords_a, sort_work_view(sc.sort_work_p, P).hist, s);
    SortDyn dd;
    dd.bits_dev = sc.ctrl + CTRL_DEPTH_BITS;
    dd.max_passes = DEPTH_MAX_PASSES;
    sort_pairs(sc.dwords_a, nullptr, sc.dwords_b, sc.perm, sc.dwords_a, nullptr, P, dplan, true, sc.sort_work_p, s, &dd);
    LSR_STAGE("depth_order");
  }
  prof(g_ev_f, 3, s);
  // the sorted words always end up in keys_b: the unsorted ones start in keys_a (odd pass count) or keys_b (even)
  uint64_t* kin = (plan.passes & 1) ? sc.keys_a : sc.keys_b;
  // LSR_LEGACY_BINNING=1 (synchronous forward only): the round-1 kernels — stand-alone scan, slot-walking duplication,
  // separate range kernel — kept selectable for A/B measurements and for bisecting a parity failure.
  static const bool legacy_env = getenv("LSR_LEGACY_BINNING") != nullptr;
  if (legacy_env && sync) {
    const int64_t R = bound;
    LSR_CUDA(cudaMemsetAsync(img.ranges, 0, (size_t)T * sizeof(uint2), s));  // rasterizer_impl.cu:313
    if (R > 0) launch_scan_tiles(P, g.tiles, sc.perm, g.offsets, sc.scan_status, sc.ctrl, s);
    prof(g_ev_f, 4, s);
    if (R > 0) {
      sort_prepare(sc.sort_work, R, plan, s);
      launch_duplicate(P, gx, gy, sc.perm, radii, g, kin, list_capacity, plan, sort_work_view(sc.sort_work, R).hist, s);
    }
    prof(g_ev_f, 5, s);
    if (R > 0) sort_pairs(kin, nullptr, sc.keys_b, (uint32_t*)list, sc.keys_a, nullptr, R, plan, true, sc.sort_work, s);
    prof(g_ev_f, 6, s);
    if (R > 0) launch_tile_ranges(R, sc.keys_b, id_bits, img.ranges, s);
    launch_finalize_tiles(T, img.ranges, img.tile_order, false, sc.ctrl, list_capacity, stamp, img.status, s);
    LSR_STAGE("legacy binning");
  } else {
    prof(g_ev_f, 4, s);  // (the scan of the tile counts is part of the duplication kernel)
    if (bound > 0) sort_prepare(sc.sort_work, bound, plan, s);
    launch_dup_scan(P, gx, gy, sc.perm, radii, g, kin, list_capacity, plan, sc.scan_status, sc.ctrl,
                    sort_work_view(sc.sort_work, bound).hist, s);
    LSR_STAGE("duplicate");
    prof(g_ev_f, 5, s);
    if (bound > 0) {
      SortDyn sd;
      sd.n_dev = sc.ctrl + CTRL_NUM_RENDERED;
      sd.ranges = img.ranges;
      sd.tile_shift = id_bits;
      sort_pairs(kin, nullptr, sc.keys_b, (uint32_t*)list, sc.keys_a, nullptr, bound, plan, true, sc.sort_work, s, &sd);
      LSR_STAGE("sort");
    }
    prof(g_ev_f, 6, s);
    launch_finalize_tiles(T, img.ranges, img.tile_order, true, sc.ctrl, list_capacity, stamp, img.status, s);
    LSR_STAGE("tile_ranges");
  }
  prof(g_ev_f, 7, s);
  launch_render_fwd(W, H, img.ranges, img.tile_order, (const uint32_t*)list, g.rec, frame->bg,
                    frame->include_feature != 0, img.final_T, img.n_contrib, out_color, out_feature, s);
  LSR_STAGE("render_fwd");
  prof(g_ev_f, 8, s);
  if (!sync) {
    launch_poison_outputs(img.status, out_color, 3 * N, frame->include_feature ? out_feature : nullptr, 3 * N, s);
    if (status_host)
      LSR_CUDA(cudaMemcpyAsync(status_host, img.status, LSR_STATUS_WORDS * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    LSR_STAGE("status");
  }
  if (g_profile) g_have_f = true;
  return LSR_OK;
}

int lsr_forward(const lsr_frame* frame, const lsr_inputs* in, float* out_color, float* out_feature, int32_t* radii,
                void* geom, void* image, void* list, int64_t list_capacity, void* scratch, size_t scratch_nbytes,
                int64_t* num_rendered, void* stream) {
  return forward_impl(frame, in, out_color, out_feature, radii, geom, image, list, list_capacity, scratch,
                      scratch_nbytes, num_rendered, true, 0u, nullptr, (cudaStream_t)stream);
}

int lsr_forward_nosync(const lsr_frame* frame, const lsr_inputs* in, float* out_color, float* out_feature,
                       int32_t* radii, void* geom, void* image, void* list, int64_t list_capacity, void* scratch,
                       size_t scratch_nbytes, uint32_t stamp, uint32_t* status_host, void* stream) {
  if (frame && frame->debug) return fail(LSR_ERR_INVALID, "debug=1 synchronises after every stage: use lsr_forward");
  return forward_impl(frame, in, out_color, out_feature, radii, geom, image, list, list_capacity, scratch,
                      scratch_nbytes, nullptr, false, stamp, status_host, (cudaStream_t)stream);
}

int lsr_forward_status(int32_t width, int32_t height, const void* image, uint32_t* status, void* stream) {
  if (!image || !status || width <= 0 || height <= 0) return fail(LSR_ERR_INVALID, "bad argument");
  ImageView img = image_view(const_cast<void*>(image), width, height);
  LSR_CUDA(cudaMemcpyAsync(status, img.status, LSR_STATUS_WORDS * sizeof(uint32_t), cudaMemcpyDeviceToHost,
                           (cudaStream_t)stream));
  LSR_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return LSR_OK;
}

int lsr_forward_feature_level(int32_t P, int32_t width, int32_t height, const int32_t* radii, void* geom,
                              const void* image, const void* list, const float* language_feature, float* out_feature,
                              void* stream) {
  if (P < 0 || width < 0 || height < 0) return fail(LSR_ERR_INVALID, "negative size");
  if (!image || !out_feature) return fail(LSR_ERR_INVALID, "null image / out_feature");
  if (P > 0 && (!radii || !geom || !language_feature)) return fail(LSR_ERR_INVALID, "null radii / geom / feature");
  cudaStream_t s = (cudaStream_t)stream;
  if (P == 0) {  // as lsr_forward: an empty scene gives zero images (rasterize_points.cu:91)
    LSR_CUDA(cudaMemsetAsync(out_feature, 0, 3 * (size_t)width * height * sizeof(float), s));
    return LSR_OK;
  }
  GeomView g = geom_view(geom, P);
  ImageView img = image_view(const_cast<void*>(image), width, height);
  launch_refeature_render(P, width, height, radii, language_feature, g.rec, img.ranges,
                          img.tile_order, (const uint32_t*)list,
                          out_feature, s);
  LSR_CUDA(cudaGetLastError());
  return LSR_OK;
}

int lsr_backward(const lsr_frame* frame, const lsr_inputs* in, const int32_t* radii, const void* geom,
                 const void* image, const void* list, const float* dL_dcolor, const float* dL_dfeature,
                 uint32_t need_mask, const lsr_grads* grads, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  int rc = check_frame(frame, in);
  if (rc != LSR_OK) return rc;
  if (!grads) return fail(LSR_ERR_INVALID, "null grads");
  const int P = frame->P, W = frame->width, H = frame->height;
  if (P == 0) return LSR_OK;
  if (!radii || !geom || !image) return fail(LSR_ERR_INVALID, "null radii / geom / image");
  const lsr_grads& gr = *grads;
  const bool feat = frame->include_feature != 0 && dL_dfeature != nullptr;
  const bool dcol = dL_dcolor != nullptr;
  const bool geom_need = (need_mask & LSR_NEED_GEOMETRY) != 0;
  if (!gr.dL_dmeans2D) return fail(LSR_ERR_INVALID, "dL_dmeans2D buffer is required");
  const bool have_acc = gr.blend_acc != nullptr && blend_acc_supported();
  if (geom_need && ((!gr.dL_dconic && !have_acc) || !gr.dL_dopacity))
    return fail(LSR_ERR_INVALID, "dL_dconic / dL_dopacity required");
  if (dcol && !gr.dL_dcolors && !have_acc)
    return fail(LSR_ERR_INVALID, "dL_dcolors buffer required when dL_dcolor is given");
  if (feat && !gr.dL_dlanguage_feature) return fail(LSR_ERR_INVALID, "dL_dlanguage_feature buffer required");
  if (((((uintptr_t)gr.dL_dconic) | ((uintptr_t)gr.dL_drotations)) & 15) != 0)
    return fail(LSR_ERR_INVALID, "dL_dconic and dL_drotations must be 16-byte aligned (read / written as float4)");
  if ((((uintptr_t)gr.dL_dcov3D) & 7) != 0)
    return fail(LSR_ERR_INVALID, "dL_dcov3D must be 8-byte aligned (written as float2)");
  if (g_profile) g_have_b = false;
  prof(g_ev_b, 0, s);
  // The packed accumulation record (lsr_grads::blend_acc): the blend kernel adds there and the kernel that follows
  // (per-Gaussian backward, else unpack) writes every element of the per-tensor buffers — only the record is zeroed.
  const bool run_pre_k = geom_need || (in->shs && (gr.dL_dsh || gr.dL_dcolors_masked) && dcol);
  float* acc = have_acc ? gr.blend_acc : nullptr;
  if (acc && (((uintptr_t)acc) & 63) != 0) return fail(LSR_ERR_INVALID, "blend_acc must be 64-byte aligned");
  if (acc && run_pre_k && geom_need && !gr.dL_dopacity) return fail(LSR_ERR_INVALID, "dL_dopacity required");
  // zero what the blend kernel accumulates into; adjacent buffers are merged into one memset
  {
    std::vector<std::pair<char*, size_t>> z;
    const size_t n = (size_t)P * sizeof(float);
    auto add = [&](float* p, size_t k) { if (p) z.push_back({(char*)p, n * k}); };
    if (acc) {
      add(acc, ACC_FLOATS);
      if (!dcol) add(gr.dL_dcolors, 3);  // (nothing is accumulated for it and nobody would write it)
      if (!feat) add(gr.dL_dlanguage_feature, 3);
      add(gr.dL_dconic, 4);  // (optional with a record: nothing ever lands in it)
      if (!run_pre_k) add(gr.dL_dopacity, 1);
    } else {
      add(gr.dL_dmeans2D, 3);
      add(gr.dL_dconic, 4);
      add(gr.dL_dopacity, 1);
      add(gr.dL_dcolors, 3);
      add(gr.dL_dlanguage_feature, 3);
    }
    // outputs the per-Gaussian kernel will not write in this configuration
    const bool run_pre = geom_need || (in->shs && (gr.dL_dsh || gr.dL_dcolors_masked) && dcol);
    if (!run_pre) { add(gr.dL_dmeans3D, 3); add(gr.dL_dcov3D, 6); add(gr.dL_dscales, 3); add(gr.dL_drotations, 4); }
    if (!run_pre || !in->shs || !(gr.dL_dcolors || acc)) { if (gr.dL_dsh && frame->M > 0) add(gr.dL_dsh, (size_t)frame->M * 3); }
    if (run_pre && !in->scales) { add(gr.dL_dscales, 3); add(gr.dL_drotations, 4); }
    std::sort(z.begin(), z.end());
    size_t i = 0;
    while (i < z.size()) {
      char* b = z[i].first;
      char* e = b + z[i].second;
      size_t j = i + 1;
      while (j < z.size() && z[j].first <= e) { e = std::max(e, z[j].first + z[j].second); j++; }
      LSR_CUDA(cudaMemsetAsync(b, 0, (size_t)(e - b), s));
      i = j;
    }
  }
  prof(g_ev_b, 1, s);
  GeomView g = geom_view(const_cast<void*>(geom), P);
  ImageView img = image_view(const_cast<void*>(image), W, H);
  launch_render_bwd(W, H, img.ranges, img.tile_order, (const uint32_t*)list, g.rec, frame->bg, frame->include_feature != 0,
                    img.final_T, img.n_contrib, dL_dcolor, dL_dfeature, need_mask, gr.dL_dmeans2D, gr.dL_dconic,
                    gr.dL_dopacity, gr.dL_dcolors, gr.dL_dlanguage_feature, acc, s);
  LSR_STAGE("render_bwd");
  prof(g_ev_b, 2, s);
  if (geom_need || (in->shs && (gr.dL_dsh || gr.dL_dcolors_masked) && dcol)) {
    if (!gr.dL_dconic && !acc) return fail(LSR_ERR_INVALID, "dL_dconic required for the per-Gaussian backward");
    const float focal_y = H / (2.0f * frame->tanfovy);
    const float focal_x = W / (2.0f * frame->tanfovx);
    launch_preprocess_bwd(*frame, *in, focal_x, focal_y, radii, g, need_mask, gr, acc, s);
    LSR_STAGE("preprocess_bwd");
  } else if (acc) {
    launch_unpack_acc(P, acc, gr.dL_dmeans2D, dcol ? gr.dL_dcolors : nullptr, feat ? gr.dL_dlanguage_feature : nullptr, s);
    LSR_STAGE("unpack_acc");
  }
  prof(g_ev_b, 3, s);
  if (g_profile) g_have_b = true;
  (void)feat;
  return LSR_OK;
}

}  // extern "C"
#pragma GCC visibility pop
