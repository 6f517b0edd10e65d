/* lsr_api.h — C ABI of liblsr.so, the B200-native (sm_100a) Gaussian-splatting rasterizer.
 *
 * This is the drop-in boundary for LangSplat's hot path.  The reference has no C ABI of its own; its
 * boundary is the pybind module `diff_gaussian_rasterization._C` (RAST/ext.cpp:15-19, RAST =
 * submodules/langsplat-rasterization) whose three functions hand raw device pointers to
 * CudaRasterizer::Rasterizer::{forward,backward,markVisible} (RAST/cuda_rasterizer/rasterizer.h:20-92).
 * Each entry point below replaces one of those and cites it.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer to float32 / int32 data allocated by the caller (PyTorch),
 *    except where a parameter is documented as "host";
 *  - the library never allocates tensors, never frees caller memory, launches only on `stream`;
 *  - return value: LSR_OK (0) or a negative lsr_status; lsr_last_error() gives a thread-local message;
 *  - `stream` is a cudaStream_t passed as void* so that this header needs no CUDA headers.
 */
#ifndef LSR_API_H_INCLUDED
#define LSR_API_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LSR_VERSION 3

typedef enum lsr_status {
  LSR_OK = 0,
  LSR_ERR_INVALID = -1,     /* bad argument (null pointer, P < 0, exclusive inputs both / neither given) */
  LSR_ERR_CAPACITY = -2,    /* instance list / scratch too small: *num_rendered holds the needed count, call again */
  LSR_ERR_CUDA = -3,        /* a CUDA call failed; message in lsr_last_error() */
  LSR_ERR_PREFILTERED = -4  /* prefiltered=1 but a Gaussian was culled (reference: __trap, auxiliary.h:156-160) */
} lsr_status;

/* Per-call constants.  Replaces GaussianRasterizationSettings (RAST/diff_gaussian_rasterization/__init__.py:166-179)
 * plus the scalar arguments of RasterizeGaussiansCUDA (RAST/rasterize_points.cu:35-57). */
typedef struct lsr_frame {
  int32_t P;               /* number of Gaussians */
  int32_t M;               /* SH coefficients per Gaussian (sh.size(1)); 0 when colors are precomputed */
  int32_t D;               /* active SH degree 0..3 */
  int32_t width, height;
  float tanfovx, tanfovy;
  float scale_modifier;
  int32_t prefiltered;
  int32_t debug;           /* 1: synchronise + check after every stage (CHECK_CUDA, auxiliary.h:166-173) */
  int32_t include_feature; /* blend the 3-channel language feature too */
  const float* bg;         /* [3] */
  const float* viewmatrix; /* [16], transposed world->view, flat index m[0]*x + m[4]*y + m[8]*z + m[12] */
  const float* projmatrix; /* [16], full projection, same convention */
  const float* campos;     /* [3] */
} lsr_frame;

/* Per-Gaussian inputs (null = absent).  Exactly one of shs / colors_precomp, and exactly one of
 * (scales, rotations) / cov3D_precomp — RAST/diff_gaussian_rasterization/__init__.py:201-205. */
typedef struct lsr_inputs {
  const float* means3D;          /* [P,3] */
  const float* shs;              /* [P,M,3] */
  const float* colors_precomp;   /* [P,3] */
  const float* language_feature; /* [P,3] (required when include_feature) */
  const float* opacities;        /* [P] */
  const float* scales;           /* [P,3] */
  const float* rotations;        /* [P,4] */
  const float* cov3D_precomp;    /* [P,6] */
} lsr_inputs;

/* Gradient outputs of lsr_backward (null = not wanted).  Order/shape as returned by
 * RasterizeGaussiansBackwardCUDA (RAST/rasterize_points.cu:129-224). */
typedef struct lsr_grads {
  float* dL_dmeans3D;           /* [P,3] */
  float* dL_dmeans2D;           /* [P,3] (z component always 0) */
  float* dL_dsh;                /* [P,M,3] */
  float* dL_dcolors;            /* [P,3] */
  float* dL_dlanguage_feature;  /* [P,3] */
  float* dL_dopacity;           /* [P] */
  float* dL_dscales;            /* [P,3] */
  float* dL_drotations;         /* [P,4] */
  float* dL_dcov3D;             /* [P,6], 8-byte aligned */
  float* dL_dconic;             /* [P,4] (internal in the reference; exposed for tests) */
  float* dL_dcolors_masked;     /* [P,3] optional: the SH gradient's per-view factor (dL_dcolor with the clamp mask of
                                   backward.cu:31-34 applied, zero for culled Gaussians).  Given instead of dL_dsh it
                                   makes lsr_backward skip the 16x larger dL_dsh; see lsr_sh_gradient_from_views */
  float* blend_acc;             /* [P,16] optional scratch, 64-byte aligned: the blend backward's packed accumulation
                                   record (one 64-byte record per Gaussian, added to with 16-byte vector reductions
                                   instead of twelve scalar atomics into five tensors — the reference's atomicAdds of
                                   backward.cu:552-582).  With it dL_dconic and (when shs are given) dL_dcolors may be
                                   null, and lsr_backward zero-fills only the record; every gradient buffer passed is
                                   still fully written.  Contents after the call are unspecified. */
} lsr_grads;

/* need_mask bits for lsr_backward: which groups of gradients the caller will read
 * (ctx.needs_input_grad on the Python side).  LSR_NEED_ALL reproduces the reference. */
#define LSR_NEED_GEOMETRY 1u /* means3D, opacity, scales, rotations, cov3D (=> conic/opacity accumulation + K8/K9) */
#define LSR_NEED_COLOR 2u    /* colors_precomp or sh */
#define LSR_NEED_FEATURE 4u  /* language_feature */
#define LSR_NEED_MEANS2D 8u  /* screen-space positional gradient */
#define LSR_NEED_ALL 15u

/* ---- buffer sizes.  The caller allocates byte buffers (>=256-B aligned) of at least these sizes. ----
 * geom  : per-call saved state, P-sized   (replaces geomBuffer,    rasterizer_impl.cu:155-170)
 * image : per-call saved state, W*H-sized (replaces imgBuffer,     rasterizer_impl.cu:172-179)
 * list  : per-call saved sorted Gaussian ids, 4 B * capacity (the part of binningBuffer backward needs)
 * scratch: transient sort / scan workspace, may be shared by all calls on one stream
 *           (replaces the rest of binningBuffer + CUB temp storage, rasterizer_impl.cu:181-194). */
size_t lsr_geom_bytes(int32_t P);
size_t lsr_image_bytes(int32_t width, int32_t height);
size_t lsr_list_bytes(int64_t capacity);
size_t lsr_scratch_bytes(int32_t P, int32_t width, int32_t height, int64_t capacity);

/* Byte offsets of the arrays inside geom / image / scratch, for tests that compare stage by stage.
 * `which` is one of the LSR_ARR_* ids; returns (size_t)-1 for an unknown id. */
enum {
  LSR_ARR_GEOM_REC = 0,      /* float[P][12]: x, y, conic.x, conic.y, conic.z, opacity, r, g, b, f0, f1, f2 */
  LSR_ARR_GEOM_DEPTH = 1,    /* float[P] */
  LSR_ARR_GEOM_TILES = 2,    /* uint32[P] tiles_touched */
  LSR_ARR_GEOM_OFFSETS = 3,  /* uint32[P] inclusive scan of tiles_touched taken in DEPTH ORDER (LSR_ARR_SCR_DEPTH_ORDER):
                                offsets[i] = end of the instances of Gaussian depth_order[i] in the unsorted list */
  LSR_ARR_GEOM_COV3D = 4,    /* float[P][6] */
  LSR_ARR_GEOM_CLAMPED = 5,  /* uint8[P] bit ch set when colour channel ch was clamped */
  LSR_ARR_IMG_FINAL_T = 10,  /* float[H*W] */
  LSR_ARR_IMG_NCONTRIB = 11, /* uint32[H*W] */
  LSR_ARR_IMG_RANGES = 12,   /* uint32[T][2] */
  LSR_ARR_SCR_KEYS_UNSORTED = 20, /* uint64[capacity] instance words (tile << id_bits | index), sort ping-pong */
  LSR_ARR_SCR_DEPTH_ORDER = 21,   /* uint32[P] Gaussian indices in ascending depth order (stable) */
  LSR_ARR_SCR_KEYS_SORTED = 22    /* uint64[capacity] sorted instance words as the forward keeps them (use
                                     lsr_debug_sorted_keys for the reference's key format) */
};
size_t lsr_array_offset(int which, int32_t P, int32_t width, int32_t height, int64_t capacity);

/* Forward.  Replaces _C.rasterize_gaussians -> RasterizeGaussiansCUDA (RAST/rasterize_points.cu:35-127) ->
 * CudaRasterizer::Rasterizer::forward (RAST/cuda_rasterizer/rasterizer_impl.cu:198-355).
 * out_color [3,H,W]; out_feature [3,H,W] (ignored when !include_feature); radii int32[P].  All three are fully
 * written (no pre-zeroing needed).  *num_rendered (HOST) receives R, the number of Gaussian/tile instances.
 * If R > list_capacity returns LSR_ERR_CAPACITY having written only radii and *num_rendered. */
int lsr_forward(const lsr_frame* frame, const lsr_inputs* in, float* out_color, float* out_feature, int32_t* radii,
                void* geom, void* image, void* list, int64_t list_capacity, void* scratch, size_t scratch_bytes,
                int64_t* num_rendered, void* stream);

/* Forward without any host synchronisation (the reference reads the instance count back in the
 * middle of every frame, rasterizer_impl.cu:284).  Same device work and same results as lsr_forward, but the instance
 * count stays on the device: launch geometry comes from `list_capacity`, the kernels read the count (and the range of
 * the depth word) from device memory, and nothing is returned to the host before the call ends — so the call can be
 * captured in a CUDA graph and the host runs ahead of the GPU.
 * What the device found out is kept in the call's `image` state as LSR_STATUS_WORDS 32-bit words (lsr_forward_status
 * reads them, synchronising) and, if `status_host` is given (PINNED host memory, LSR_STATUS_WORDS words), also copied
 * there asynchronously on `stream`: the copy is complete once status_host[LSR_STATUS_STAMP] == stamp (the caller
 * picks `stamp` != 0 and clears the word first).
 * If the instance count exceeds list_capacity the frame cannot be rendered: LSR_STATUS_FLAG_OVERFLOW is raised and
 * out_color / out_feature are filled with NaN, so that an incomplete image can never pass for a result; the caller
 * repeats the frame with the capacity the status reports.  prefiltered=1 violations are reported the same way
 * (LSR_STATUS_FLAG_PREFILTERED) instead of LSR_ERR_PREFILTERED. */
enum {
  LSR_STATUS_NUM_RENDERED = 0, /* R, the number of Gaussian/tile instances of the frame (may exceed the capacity) */
  LSR_STATUS_FLAGS = 1,        /* LSR_STATUS_FLAG_* */
  LSR_STATUS_DEPTH_MIN = 2,    /* float bits of the smallest / largest view depth among the visible Gaussians */
  LSR_STATUS_DEPTH_MAX = 3,
  LSR_STATUS_CAPACITY = 4,     /* the list_capacity the frame ran with */
  LSR_STATUS_STAMP = 5,        /* the caller's stamp, written last */
  LSR_STATUS_WORDS = 8
};
enum { LSR_STATUS_FLAG_PREFILTERED = 1, LSR_STATUS_FLAG_OVERFLOW = 2 };
int lsr_forward_nosync(const lsr_frame* frame, const lsr_inputs* in, float* out_color, float* out_feature,
                       int32_t* radii, void* geom, void* image, void* list, int64_t list_capacity, void* scratch,
                       size_t scratch_bytes, uint32_t stamp, uint32_t* status_host, void* stream);
/* Status words of a forward (either kind) from its `image` state; synchronises `stream`.  status: HOST, 8 words. */
int lsr_forward_status(int32_t width, int32_t height, const void* image, uint32_t* status, void* stream);

/* Another feature level on the same geometry (LangSplat's three SAM levels share every Gaussian
 * attribute except language_feature, and render.py renders them back to back).  Takes the saved state of an
 * lsr_forward made with include_feature, replaces the language feature held in its per-Gaussian records by
 * `language_feature` [P,3] and re-blends the feature image only — no preprocess, no binning, no RGB.
 * out_feature [3,H,W] is fully written and equals what a full lsr_forward with that feature tensor would produce.
 * Afterwards the state describes the new level. */
int lsr_forward_feature_level(int32_t P, int32_t width, int32_t height, const int32_t* radii, void* geom,
                              const void* image, const void* list, const float* language_feature, float* out_feature,
                              void* stream);

/* Backward.  Replaces _C.rasterize_gaussians_backward -> RasterizeGaussiansBackwardCUDA
 * (RAST/rasterize_points.cu:129-224) -> Rasterizer::backward (rasterizer_impl.cu:359-463).
 * dL_dcolor / dL_dfeature are [3,H,W]; dL_dcolor may be null (treated as zero, as in feature-only training);
 * dL_dfeature may be null when !include_feature.  Every non-null lsr_grads pointer is fully written
 * (the library zero-fills what it accumulates into; no pre-zeroing by the caller). */
int lsr_backward(const lsr_frame* frame, const lsr_inputs* in, const int32_t* radii, const void* geom,
                 const void* image, const void* list, const float* dL_dcolor, const float* dL_dfeature,
                 uint32_t need_mask, const lsr_grads* grads, void* stream);

/* dL_dsh from per-view factors (data-parallel / multi-view steps).  One view's SH gradient is rank-1 per Gaussian:
 * dL_dsh[k][ch] = basis_k(normalize(mean - campos)) * g[ch] (RAST/cuda_rasterizer/backward.cu:36-129), with g that
 * view's dL_dcolors_masked.  This sums V views: means3D [P,3], campos [V,3], g [V,P,3] -> dL_dsh [P,M,3] (fully
 * written; rows above the active degree D are zero).  A view-sharded step exchanges the 12-byte factors (all-gather)
 * instead of all-reducing the 192-byte products. */
int lsr_sh_gradient_from_views(int32_t P, int32_t M, int32_t D, int32_t n_views, const float* means3D,
                               const float* campos, const float* g, float* dL_dsh, void* stream);

/* The same sum with every view's factor and camera position given by its own device pointer: g_views[v] -> [P,3],
 * campos_views[v] -> 3 floats (both arrays live on the host, n_views entries).  The pointers may be peer memory mapped
 * with lsr_peer_open: the kernel then reads the other ranks' factors in place over NVLink and no all-gather is needed.
 * More than LSR_MAX_VIEWS views are processed in chunks. */
#define LSR_MAX_VIEWS 64
int lsr_sh_gradient_from_view_ptrs(int32_t P, int32_t M, int32_t D, int32_t n_views, const float* means3D,
                                   const float* const* g_views, const float* const* campos_views, float* dL_dsh,
                                   void* stream);

/* ---- peer memory for the view-sharded step (one process per GPU of one node; no reference counterpart: the
 * reference is single-GPU, train.py:84-92) ----
 * lsr_peer_alloc: device buffer of `bytes` (zero-filled) that other processes can map, and its 64-byte handle (a CUDA
 *   IPC handle) to hand to them by any channel (torch.distributed all_gather in langsplat_b200.dist).
 * lsr_peer_open / lsr_peer_close: map / unmap another process's buffer; lsr_peer_free: release an own buffer.
 * The first LSR_PEER_HEADER_BYTES of every buffer belong to the library (barrier words); payload starts after them.
 * lsr_peer_barrier: stream-ordered barrier across `world` ranks; headers[r] = rank r's buffer as seen from this
 *   process (own pointer at headers[rank]).  `epoch` must grow by one per call on every rank.  Work enqueued after it
 *   sees everything the peers enqueued before theirs.  A peer that does not arrive within the time-out (lsr_peer_set_timeout_ms) is reported by
 *   lsr_peer_status (returns LSR_ERR_CUDA, "peer barrier timed out") instead of hanging the device.
 * lsr_peer_signal / lsr_peer_wait: the two halves of the barrier, for a rank that has local work to run between
 *   "my remote writes are done" and "everybody's are".
 * lsr_peer_sync_auto: the same barrier (mode 3), signal (1) or wait (2) with the epoch counted on the device (in the
 *   own header: a signal advances it, a wait uses the value its signal left) instead of passed by the host, so that
 *   the launch is identical every step and a step with its exchange can be replayed from a CUDA graph.  Every rank
 *   must issue the same sequence of calls; do not mix with the explicit-epoch calls on the same buffers.
 * lsr_peer_allreduce: two-shot sum over the ranks through peer memory.  in[r] / out[r] = rank r's input / output
 *   vector of n_floats floats (16-byte aligned, n_floats % 4 == 0) as mapped in this process.  This rank reduces its
 *   own 1/world chunk from all inputs (fixed rank order: every rank ends with the same bits) and stores it into every
 *   rank's output.  Order it with the barrier: inputs complete on all ranks before it (lsr_peer_barrier), outputs
 *   complete once every rank has signalled after it (lsr_peer_signal ... lsr_peer_wait). */
#define LSR_PEER_HANDLE_BYTES 64
#define LSR_PEER_HEADER_BYTES 256
#define LSR_MAX_PEERS 16
int lsr_peer_alloc(size_t bytes, void** ptr, uint8_t* handle);
int lsr_peer_open(const uint8_t* handle, void** ptr);
int lsr_peer_close(void* ptr);
int lsr_peer_free(void* ptr);
int lsr_peer_barrier(int32_t rank, int32_t world, void* const* headers, uint32_t epoch, void* stream);
int lsr_peer_signal(int32_t rank, int32_t world, void* const* headers, uint32_t epoch, void* stream);
int lsr_peer_wait(int32_t rank, int32_t world, void* const* headers, uint32_t epoch, void* stream);
int lsr_peer_sync_auto(int32_t rank, int32_t world, void* const* headers, int32_t mode, void* stream);
int lsr_peer_allreduce(int32_t rank, int32_t world, const float* const* in, float* const* out, int64_t n_floats,
                       void* stream);
int lsr_peer_status(const void* own_buffer, void* stream);
/* Time-out of the barrier's wait (default 120 000 ms; process-wide).  A rank that does not arrive in time raises a
 * sticky flag in the waiting rank's header: lsr_peer_status then fails, and lsr_peer_poison_on_timeout — to be
 * enqueued behind every kernel that consumed peer data — overwrites `data` with NaN so that an incomplete sum can
 * never pass for a gradient (one word read per thread when no time-out happened). */
int lsr_peer_set_timeout_ms(uint64_t ms);
int lsr_peer_poison_on_timeout(const void* own_buffer, float* data, int64_t n_floats, void* stream);

/* Replaces _C.mark_visible -> markVisible (rasterize_points.cu:226-245) -> checkFrustum (rasterizer_impl.cu:54-66). */
int lsr_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream);

/* Fused (masked) L1 loss on the rendered images — the consumer right after the rasterizer in the training step
 * (SURVEY.md 8(f) row 1).  Replaces l1_loss(pred * mask, gt * mask) = torch.abs(pred*mask - gt*mask).mean()
 * (utils/loss_utils.py:17-18 as called from train.py:96-99).  pred, gt: [C,N] (N = H*W); mask: [N] broadcast over
 * the channels, or null (= ones, the RGB call of train.py:102).  *loss (DEVICE scalar) receives the mean. */
int lsr_l1_loss_forward(const float* pred, const float* gt, const float* mask, int32_t C, int64_t N, float* loss,
                        void* stream);
/* dL_dpred[C,N] = (*grad_out) * sign(pred*mask - gt*mask) * mask / (C*N); grad_out is a DEVICE scalar. */
int lsr_l1_loss_backward(const float* pred, const float* gt, const float* mask, int32_t C, int64_t N,
                         const float* grad_out, float* dL_dpred, void* stream);
/* The same two with the ground truth as the 8-bit image it is stored as: gt = u / 255 (IEEE division, the value
 * PILtoTorch, utils/general_utils.py:21-28, forms on the host before scene/cameras.py:39 uploads it as float32) is
 * computed inside the kernels, so the float copy of the image never exists in device memory. */
int lsr_l1_loss_forward_u8(const float* pred, const uint8_t* gt, const float* mask, int32_t C, int64_t N, float* loss,
                           void* stream);
int lsr_l1_loss_backward_u8(const float* pred, const uint8_t* gt, const float* mask, int32_t C, int64_t N,
                            const float* grad_out, float* dL_dpred, void* stream);

/* Fused per-Gaussian activations in front of the rasterizer: scales = exp(scaling)
 * (scene/gaussian_model.py:33,134-136), rotations = normalize(rotation) = q / max(||q||, 1e-12) (:40,138-140),
 * opacities = sigmoid(opacity) (:37,153-155), language feature f / (||f|| + 1e-9)
 * (gaussian_renderer/__init__.py:86-88).  scaling [P,3], rotation [P,4] (16-byte aligned), opacity [P,1],
 * feature [P,3]; an input given as null is skipped together with its output. */
int lsr_activate_forward(int32_t P, const float* scaling, const float* rotation, const float* opacity,
                         const float* feature, float* scales_out, float* rotations_out, float* opacities_out,
                         float* feature_out, void* stream);
/* Gradients w.r.t. the raw parameters from the gradients w.r.t. the activated ones (recomputed from the raw inputs,
 * nothing saved by the forward).  A quantity is processed when its d_* output is non-null; its raw input and its
 * upstream gradient g_* must then be given. */
int lsr_activate_backward(int32_t P, const float* scaling, const float* rotation, const float* opacity,
                          const float* feature, const float* g_scales, const float* g_rotations,
                          const float* g_opacities, const float* g_feature, float* d_scaling, float* d_rotation,
                          float* d_opacity, float* d_feature, void* stream);

/* Ground-truth language-feature map of one camera and level (replaces the CPU gather of
 * Camera.get_language_feature, scene/cameras.py:58-92).  seg: one level of the seg map, H rows of W values
 * (float32 as preprocess.py:120,173 saves it, or int32), consecutive rows seg_row_stride elements apart (a larger map
 * is cropped to its top-left H x W like the reference's meshgrid does); table: feature_map [K,C].
 * out [C,H,W] = table[seg] (index -1 selects the last row, as the reference's negative index does),
 * mask_u8 / mask_f [H,W] = (seg != -1), either may be null.  *bad (device int, may be null; caller zeroes it) counts
 * indices outside [-K, K), for which PyTorch would raise; those pixels get zeros. */
int lsr_language_feature_map(const void* seg, int32_t seg_is_float, int64_t seg_row_stride, int32_t width,
                             int32_t height, const float* table, int32_t K, int32_t C, float* out, uint8_t* mask_u8,
                             float* mask_f, int32_t* bad, void* stream);

/* Fused Adam step over all parameter groups in one launch (replaces optimizer.step() of
 * torch.optim.Adam(l, lr=0.0, eps=1e-15), scene/gaussian_model.py:198-227, train.py:136-138).  PyTorch's single-tensor
 * Adam arithmetic (amsgrad off, no weight decay); `step` is the 1-based step count AFTER the increment (PyTorch's
 * state['step']).  Each tensor: param / exp_avg / exp_avg_sq updated in place from grad, `numel` floats, its own lr;
 * row_width = floats per Gaussian (used only with `visible`).  visible [P] uint8 or null: when given, only rows of
 * visible Gaussians are updated, moments included ("sparse Adam"); null = the reference's dense step. */
#define LSR_ADAM_MAX_TENSORS 8
typedef struct lsr_adam_tensor {
  float* param;
  const float* grad;
  float* exp_avg;
  float* exp_avg_sq;
  int64_t numel;
  int32_t row_width;
  float lr;
} lsr_adam_tensor;
int lsr_adam_step(int32_t n_tensors, const lsr_adam_tensor* tensors, int64_t step, double beta1, double beta2,
                  double eps, const uint8_t* visible, void* stream);

/* Eval-side decode of rendered language features: the autoencoder's decoder MLP followed by
 * the CLIP relevancy of every pixel — replaces Autoencoder.decode (autoencoder/model.py:42-46, layers built at :18-25;
 * call site eval/evaluate_iou_loc.py:258-261) and OpenCLIPNetwork.get_relevancy / get_max_across
 * (eval/openclip_encoder.py:42-56, 96-111) with one fused kernel.
 * lsr_decoder: n_layers Linear layers, layer l maps dims[l] -> dims[l+1] with nn.Linear's weight [dims[l+1], dims[l]]
 * (row-major) and bias [dims[l+1]]; ReLU between layers, none after the last; the result is divided by its L2 norm.
 * packed_tf32 / packed_f16 (device, lsr_decoder_packed_bytes(dec, precision) bytes, 16-byte aligned) are the weight
 * blobs lsr_decoder_pack writes for the tensor-core variants (pack once per checkpoint); each may be null when its
 * variant is not used.
 * lsr_decode: in [n, dims[0]] (a flattened [H,W,3] feature image) ->
 *   out [n, dims[L]] unit vectors (may be null), and/or
 *   relevancy [n_pos, n]: for positive phrase j the first column of get_relevancy(embed, j), i.e.
 *   min over negatives k of softmax(10 * (sim_j, sim_neg_k))[0] (may be null; then phrases may be null too).
 * phrases [n_pos + n_neg, dims[L]]: unit phrase embeddings, positives first (pos_embeds then neg_embeds);
 * n_pos + n_neg <= LSR_DECODER_MAX_PHRASES, n_neg >= 1 when relevancy is requested.
 * precision: LSR_DECODE_FP32 = IEEE fp32 on the FP32 pipe (the reference's arithmetic up to summation order);
 * LSR_DECODE_TF32 / LSR_DECODE_F16 = tcgen05 tensor cores with fp32 accumulation in TMEM, operands rounded to the
 * nearest TF32 (10 mantissa bits) or fp16 (10 mantissa bits, clamped to +-65504) value; both need their packed blob,
 * layer widths that are multiples of 16 and at most 512 (256 for all but the last), inputs at most 256 wide;
 * otherwise LSR_ERR_INVALID.  work: device scratch of lsr_decode_work_bytes() bytes, 16-byte aligned, private to the
 * call until it has finished on `stream` (the tensor-core variants fold the phrases into the last layer there:
 * sim_j = h . (P W)_j + P_j . b, one more N = 16 MMA pass); may be null without relevancy or with LSR_DECODE_FP32. */
#define LSR_DECODER_MAX_LAYERS 8
#define LSR_DECODER_MAX_PHRASES 16
#define LSR_DECODE_FP32 0
#define LSR_DECODE_TF32 1
#define LSR_DECODE_F16 2
#define LSR_DECODE_BF16X2 3 /* tensor cores at the reference's arithmetic: every fp32 operand carried as two bf16 numbers
                               (hi + lo, 16 mantissa bits), every product as three bf16 MMAs with fp32 accumulation */
typedef struct lsr_decoder {
  int32_t n_layers;
  int32_t dims[LSR_DECODER_MAX_LAYERS + 1];
  const float* weight[LSR_DECODER_MAX_LAYERS];
  const float* bias[LSR_DECODER_MAX_LAYERS];
  const void* packed_tf32;
  const void* packed_f16;
  const void* packed_bf16x2;
} lsr_decoder;
size_t lsr_decoder_packed_bytes(const lsr_decoder* dec, int32_t precision);
int lsr_decoder_pack(const lsr_decoder* dec, int32_t precision, void* packed, void* stream);
size_t lsr_decode_work_bytes(void);
int lsr_decode(const lsr_decoder* dec, const float* in, int64_t n, int32_t precision, float* out, const float* phrases,
               int32_t n_pos, int32_t n_neg, float* relevancy, void* work, size_t work_bytes, void* stream);
/* Profiling aid: later tensor-core lsr_decode launches make CTA 0 write clock64() stamps of its second tile into
 * device_buffer (>= 64 int64: tile start, then per pass "MMAs issued", "accumulator complete", "epilogue done");
 * null switches it off. */
int lsr_debug_decode_trace(void* device_buffer);

/* Stable ascending radix sort of (uint64 key, uint32 value) pairs on bits [0, end_bit) — the sort the forward uses
 * in place of cub::DeviceRadixSort::SortPairs (rasterizer_impl.cu:306-311); exported for unit tests.
 * keys_tmp / vals_tmp are n-sized ping-pong buffers; result lands in keys_out / vals_out.
 * `work` needs lsr_sort_work_bytes(n) bytes. */
size_t lsr_sort_work_bytes(int64_t n);
int lsr_sort_pairs(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                   uint64_t* keys_tmp, uint32_t* vals_tmp, int64_t n, int end_bit, void* work, size_t work_bytes,
                   void* stream);

/* lsr_sort_pairs without any allocation by caller or library.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving
 * loops, harnesses that forbid it).  Same kernels and same result as lsr_sort_pairs, but the work area and the two
 * ping-pong buffers are a static device array inside the library (14.8 MB per GPU that loads liblsr.so, held until
 * the process exits) instead of caller-provided memory: the caller passes only inputs and outputs.  The call makes
 * no cudaMalloc / cudaFree / allocating library call.
 * Limit: n <= LSR_SORT_STATIC_MAX_N pairs; a larger n is rejected with LSR_ERR_CAPACITY and a message.
 * Every call on a device shares the one array: calls that overlap on the GPU (two streams, two host threads) race —
 * issue them on one stream, or order them with events.  Nothing needs clearing between calls (the library re-zeroes
 * the counters it needs). */
#define LSR_SORT_STATIC_MAX_N 1048576
int lsr_sort_pairs_static(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                          int64_t n, int end_bit, void* stream);

/* Tests / debugging: writes the sorted instance keys of the most recent lsr_forward issued from this thread, in the
 * reference's format (tile_id << 32 | float bits of depth; point_list_keys of rasterizer_impl.cu:181-194), into
 * keys_out[num_rendered].  The forward orders the Gaussians by depth first and then sorts instance words
 * (tile_id << id_bits | Gaussian index) on the tile bits only; the reference key of an instance is its tile id with
 * the depth bits of its Gaussian.  `scratch` / `geom` are the buffers that forward was given. */
int lsr_debug_sorted_keys(const void* scratch, const void* geom, int32_t P, int64_t capacity, int64_t num_rendered,
                          uint64_t* keys_out, void* stream);

/* Per-stage device timing (CUDA events on `stream`), used by bench.py for the live roofline figures.
 * lsr_profile_enable(1) makes forward/backward record events; lsr_profile_read() synchronises the events of the
 * most recent forward+backward and writes up to n stage durations in ms, returning the number of stages.
 * Stage order: see lsr_profile_stage_name(). */
void lsr_profile_enable(int on);
int lsr_profile_read(float* ms, int n);
const char* lsr_profile_stage_name(int i);

/* Number of kernel launches issued by this library since load (for bench.py's gpu_launches). */
uint64_t lsr_launch_count(void);

const char* lsr_last_error(void);
int lsr_version(void);
/* 1 when lsr_backward uses lsr_grads::blend_acc (the default two-phase blend backward); 0 when the one-phase kernel
 * was selected (environment LSR_BWD_V1=1) — the caller must then pass dL_dconic / dL_dcolors as before. */
int lsr_blend_acc_supported(void);

#ifdef __cplusplus
}
#endif
#endif /* LSR_API_H_INCLUDED */
