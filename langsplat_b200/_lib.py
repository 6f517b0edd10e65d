"""ctypes binding of liblsr.so (include/lsr_api.h).

There is no fallback: if the CUDA library is missing or does not export the expected symbols the
import fails loudly (build it with `python -c "import __graft_entry__ as g; g.build()"` or
`make -C langsplat_b200/csrc`).
"""
from __future__ import annotations

import ctypes
import os
from ctypes import (POINTER, Structure, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_size_t, c_uint32,
                    c_uint64, c_void_p)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liblsr.so")

LSR_OK = 0
LSR_ERR_INVALID = -1
LSR_ERR_CAPACITY = -2
LSR_ERR_CUDA = -3
LSR_ERR_PREFILTERED = -4

# status words of a forward (include/lsr_api.h LSR_STATUS_*)
STATUS_NUM_RENDERED, STATUS_FLAGS, STATUS_DEPTH_MIN, STATUS_DEPTH_MAX, STATUS_CAPACITY, STATUS_STAMP, STATUS_WORDS = (
    0, 1, 2, 3, 4, 5, 8)
STATUS_FLAG_PREFILTERED, STATUS_FLAG_OVERFLOW = 1, 2

NEED_GEOMETRY, NEED_COLOR, NEED_FEATURE, NEED_MEANS2D, NEED_ALL = 1, 2, 4, 8, 15

# lsr_array_offset ids
ARR_GEOM_REC, ARR_GEOM_DEPTH, ARR_GEOM_TILES, ARR_GEOM_OFFSETS, ARR_GEOM_COV3D, ARR_GEOM_CLAMPED = 0, 1, 2, 3, 4, 5
ARR_IMG_FINAL_T, ARR_IMG_NCONTRIB, ARR_IMG_RANGES = 10, 11, 12
ARR_SCR_KEYS_UNSORTED, ARR_SCR_DEPTH_ORDER, ARR_SCR_KEYS_SORTED = 20, 21, 22


class Frame(Structure):
    _fields_ = [("P", c_int32), ("M", c_int32), ("D", c_int32), ("width", c_int32), ("height", c_int32),
                ("tanfovx", c_float), ("tanfovy", c_float), ("scale_modifier", c_float), ("prefiltered", c_int32),
                ("debug", c_int32), ("include_feature", c_int32), ("bg", c_void_p), ("viewmatrix", c_void_p),
                ("projmatrix", c_void_p), ("campos", c_void_p)]


class Inputs(Structure):
    _fields_ = [(n, c_void_p) for n in ("means3D", "shs", "colors_precomp", "language_feature", "opacities",
                                         "scales", "rotations", "cov3D_precomp")]


class Grads(Structure):
    _fields_ = [(n, c_void_p) for n in ("dL_dmeans3D", "dL_dmeans2D", "dL_dsh", "dL_dcolors",
                                         "dL_dlanguage_feature", "dL_dopacity", "dL_dscales", "dL_drotations",
                                         "dL_dcov3D", "dL_dconic", "dL_dcolors_masked", "blend_acc")]


SORT_STATIC_MAX_N = 1 << 20  # LSR_SORT_STATIC_MAX_N: pairs the allocation-free sort's static scratch holds
ABI_VERSION = 3  # LSR_VERSION of include/lsr_api.h this module's structures mirror


PEER_HANDLE_BYTES, PEER_HEADER_BYTES, MAX_PEERS, MAX_VIEWS = 64, 256, 16, 64  # include/lsr_api.h

class AdamTensor(ctypes.Structure):  # lsr_adam_tensor
    _fields_ = [("param", c_void_p), ("grad", c_void_p), ("exp_avg", c_void_p), ("exp_avg_sq", c_void_p),
                ("numel", c_int64), ("row_width", c_int32), ("lr", ctypes.c_float)]


ADAM_MAX_TENSORS = 8
DECODER_MAX_LAYERS, DECODER_MAX_PHRASES, DECODE_FP32, DECODE_TF32, DECODE_F16, DECODE_BF16X2 = 8, 16, 0, 1, 2, 3


class Decoder(ctypes.Structure):  # lsr_decoder
    _fields_ = [("n_layers", c_int32), ("dims", c_int32 * (DECODER_MAX_LAYERS + 1)),
                ("weight", c_void_p * DECODER_MAX_LAYERS), ("bias", c_void_p * DECODER_MAX_LAYERS),
                ("packed_tf32", c_void_p), ("packed_f16", c_void_p), ("packed_bf16x2", c_void_p)]


EXPORTS = {
    "lsr_geom_bytes": (c_size_t, [c_int32]),
    "lsr_image_bytes": (c_size_t, [c_int32, c_int32]),
    "lsr_list_bytes": (c_size_t, [c_int64]),
    "lsr_scratch_bytes": (c_size_t, [c_int32, c_int32, c_int32, c_int64]),
    "lsr_array_offset": (c_size_t, [c_int, c_int32, c_int32, c_int32, c_int64]),
    "lsr_forward": (c_int, [POINTER(Frame), POINTER(Inputs), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                            c_void_p, c_int64, c_void_p, c_size_t, POINTER(c_int64), c_void_p]),
    "lsr_forward_nosync": (c_int, [POINTER(Frame), POINTER(Inputs), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_int64, c_void_p, c_size_t, c_uint32, c_void_p, c_void_p]),
    "lsr_forward_status": (c_int, [c_int32, c_int32, c_void_p, c_void_p, c_void_p]),
    "lsr_forward_feature_level": (c_int, [c_int32, c_int32, c_int32, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                          c_void_p, c_void_p]),
    "lsr_backward": (c_int, [POINTER(Frame), POINTER(Inputs), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                             c_void_p, c_uint32, POINTER(Grads), c_void_p]),
    "lsr_sh_gradient_from_views": (c_int, [c_int32, c_int32, c_int32, c_int32, c_void_p, c_void_p, c_void_p, c_void_p,
                                           c_void_p]),
    "lsr_sh_gradient_from_view_ptrs": (c_int, [c_int32, c_int32, c_int32, c_int32, c_void_p, POINTER(c_void_p),
                                               POINTER(c_void_p), c_void_p, c_void_p]),
    "lsr_peer_alloc": (c_int, [c_size_t, POINTER(c_void_p), c_void_p]),
    "lsr_peer_open": (c_int, [c_void_p, POINTER(c_void_p)]),
    "lsr_peer_close": (c_int, [c_void_p]),
    "lsr_peer_free": (c_int, [c_void_p]),
    "lsr_peer_barrier": (c_int, [c_int32, c_int32, POINTER(c_void_p), c_uint32, c_void_p]),
    "lsr_peer_signal": (c_int, [c_int32, c_int32, POINTER(c_void_p), c_uint32, c_void_p]),
    "lsr_peer_wait": (c_int, [c_int32, c_int32, POINTER(c_void_p), c_uint32, c_void_p]),
    "lsr_peer_sync_auto": (c_int, [c_int32, c_int32, POINTER(c_void_p), c_int32, c_void_p]),
    "lsr_peer_allreduce": (c_int, [c_int32, c_int32, POINTER(c_void_p), POINTER(c_void_p), c_int64, c_void_p]),
    "lsr_peer_status": (c_int, [c_void_p, c_void_p]),
    "lsr_peer_set_timeout_ms": (c_int, [c_uint64]),
    "lsr_peer_poison_on_timeout": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "lsr_activate_forward": (c_int, [c_int32] + [c_void_p] * 9),
    "lsr_activate_backward": (c_int, [c_int32] + [c_void_p] * 13),
    "lsr_language_feature_map": (c_int, [c_void_p, c_int32, c_int64, c_int32, c_int32, c_void_p, c_int32, c_int32,
                                         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "lsr_adam_step": (c_int, [c_int32, c_void_p, c_int64, c_double, c_double, c_double, c_void_p, c_void_p]),
    "lsr_decoder_packed_bytes": (c_size_t, [POINTER(Decoder), c_int32]),
    "lsr_decoder_pack": (c_int, [POINTER(Decoder), c_int32, c_void_p, c_void_p]),
    "lsr_decode_work_bytes": (c_size_t, []),
    "lsr_decode": (c_int, [POINTER(Decoder), c_void_p, c_int64, c_int32, c_void_p, c_void_p, c_int32, c_int32, c_void_p,
                           c_void_p, c_size_t, c_void_p]),
    "lsr_debug_decode_trace": (c_int, [c_void_p]),
    "lsr_mark_visible": (c_int, [c_int32, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "lsr_l1_loss_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_int32, c_int64, c_void_p, c_void_p]),
    "lsr_l1_loss_backward": (c_int, [c_void_p, c_void_p, c_void_p, c_int32, c_int64, c_void_p, c_void_p, c_void_p]),
    "lsr_l1_loss_forward_u8": (c_int, [c_void_p, c_void_p, c_void_p, c_int32, c_int64, c_void_p, c_void_p]),
    "lsr_l1_loss_backward_u8": (c_int, [c_void_p, c_void_p, c_void_p, c_int32, c_int64, c_void_p, c_void_p, c_void_p]),
    "lsr_sort_work_bytes": (c_size_t, [c_int64]),
    "lsr_sort_pairs": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p,
                               c_size_t, c_void_p]),
    "lsr_debug_sorted_keys": (c_int, [c_void_p, c_void_p, c_int32, c_int64, c_int64, c_void_p, c_void_p]),
    "lsr_profile_enable": (None, [c_int]),
    "lsr_profile_read": (c_int, [POINTER(c_float), c_int]),
    "lsr_profile_stage_name": (c_char_p, [c_int]),
    "lsr_launch_count": (c_uint64, []),
    "lsr_last_error": (c_char_p, []),
    "lsr_sort_pairs_static": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "lsr_version": (c_int, []),
    "lsr_blend_acc_supported": (c_int, []),
}

_lib = None


def load() -> ctypes.CDLL:
    """Load liblsr.so and bind every symbol include/lsr_api.h declares.  Raises if anything is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: the CUDA extension is not built and langsplat_b200 has no CPU fallback. "
            "Run `make -C langsplat_b200/csrc` (needs nvcc).")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in EXPORTS.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise ImportError(f"liblsr.so does not export {name}") from e
        fn.restype = res
        fn.argtypes = args
    if lib.lsr_version() != ABI_VERSION:  # a stale build would read this module's structures with another layout
        raise ImportError(f"{LIB_PATH} has ABI version {lib.lsr_version()}, this package needs {ABI_VERSION}: "
                          "rebuild it (`make -C langsplat_b200/csrc`)")
    _lib = lib
    return lib


def last_error() -> str:
    return load().lsr_last_error().decode("utf-8", "replace")
