"""Host-side mirror of the reference's `diff_gaussian_rasterization` module over the liblsr.so C ABI.

Same names, argument meaning and error behaviour as
submodules/langsplat-rasterization/diff_gaussian_rasterization/__init__.py (RAST/...):
`GaussianRasterizationSettings` (:166-179), `GaussianRasterizer` (:181-233), `rasterize_gaussians` (:21-44) and the
autograd function (:46-164).  PyTorch is used for device memory, streams and autograd plumbing only; every kernel
lives in liblsr.so.  There is no CPU path: tensors must be CUDA float32.
"""
from __future__ import annotations

import collections
import contextlib
import ctypes
import os
from typing import Dict, NamedTuple, Optional, Tuple

import torch
import torch.nn as nn

from . import _lib as L


class GaussianRasterizationSettings(NamedTuple):  # RAST/diff_gaussian_rasterization/__init__.py:166-179
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    include_feature: bool


def cpu_deep_copy_tuple(input_tuple):  # RAST/diff_gaussian_rasterization/__init__.py:17-19
    return tuple(item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple)


class LsrError(RuntimeError):
    pass


_MSG_COLORS = "Please provide excatly one of either SHs or precomputed colors!"
_MSG_COV = "Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!"

# Per (device, stream) transient workspace and the last instance count, used as the next capacity guess.
_scratch: Dict[Tuple[int, int], torch.Tensor] = {}
_last_R: Dict[int, int] = {}
# first guess of the instance-list capacity: max(1.25 x the previous frame's count, 4 per Gaussian, 64 Ki)
_CAPACITY_PER_GAUSSIAN = 4
_INITIAL_CAPACITY_FLOOR = 1 << 16
_capacity_retries = 0  # how often lsr_forward had to be repeated with a larger list (diagnostics / tests)


# ---- forward without host synchronisation -----------------------------------------------------------------------------
# The reference reads the instance count back in the middle of every forward (rasterizer_impl.cu:284), and so does the
# default path here (lsr_forward: it has to report the count and to refuse a list that is too small before anything is
# blended).  With `set_async_forward(True)` (or LSR_ASYNC_FORWARD=1) a forward whose device has seen a frame before
# goes through lsr_forward_nosync instead: the instance list is sized from the recent frames' counts times
# `capacity_margin`, nothing is read back, the host runs ahead of the GPU and the call can be captured in a CUDA graph.
# What the device found out arrives later through pinned memory and is looked at — without waiting — at the next
# forward / backward on that device (or, waiting, by `check_async_status()`): an instance list that turned out too
# small raises LsrError there, and that frame's images were NaN-filled on the device, so an incomplete render can
# never pass for a result.  Same kernels, same results as the synchronous path.
# packed accumulation record in the backward: 0 = never (A/B), 1 = when the per-Gaussian kernel follows, 2 = always
_USE_BLEND_ACC = int(os.environ.get("LSR_BLEND_ACC", "1"))


class _AsyncConfig:
    enabled = os.environ.get("LSR_ASYNC_FORWARD", "0") not in ("", "0")
    capacity_margin = 1.5


_recent_R: Dict[int, "collections.deque"] = {}  # per device: instance counts of the last frames
_pending: Dict[int, list] = {}                  # per device: forwards whose status words have not been looked at
_status_pool: list = []                         # free pinned status buffers: (tensor, numpy view)
_stamp = [0]


def set_async_forward(enabled: bool = True, capacity_margin: Optional[float] = None) -> None:
    """Switches the forward between the reference's behaviour (one host read-back per frame) and the asynchronous
    path described above.  `capacity_margin`: head room of the instance list over the largest recent frame."""
    _AsyncConfig.enabled = bool(enabled)
    if capacity_margin is not None:
        if not capacity_margin >= 1.0:
            raise ValueError("capacity_margin must be >= 1")
        _AsyncConfig.capacity_margin = float(capacity_margin)


@contextlib.contextmanager
def async_forward(enabled: bool = True, capacity_margin: Optional[float] = None):
    saved = (_AsyncConfig.enabled, _AsyncConfig.capacity_margin)
    set_async_forward(enabled, capacity_margin)
    try:
        yield
    finally:
        _AsyncConfig.enabled, _AsyncConfig.capacity_margin = saved


def reserve_instances(device, count: int) -> None:
    """Tells the asynchronous path how many Gaussian/tile instances to expect on `device` (e.g. the maximum over the
    training cameras, found with one synchronous forward each), so that no frame outgrows its list."""
    dev = torch.device(device)
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    _recent_R.setdefault(idx, collections.deque(maxlen=16)).append(int(count))
    _last_R[idx] = max(_last_R.get(idx, 0), int(count))


def _note_R(dev_idx: int, R: int) -> None:
    _recent_R.setdefault(dev_idx, collections.deque(maxlen=16)).append(R)
    _last_R[dev_idx] = R


_status_blocks: list = []  # the pinned allocations behind the pool; never freed (a captured CUDA graph may hold them)


def _take_status(dev_idx: int):
    """A pinned status buffer.  The pool is allocated once (256 buffers, 8 KB): a pinned allocation synchronises the
    device, so there must be none once frames are in flight (nor inside a CUDA-graph capture).  A host that runs
    further ahead than the pool is deep waits for its oldest outstanding frame instead."""
    if not _status_blocks:
        block = torch.zeros((256, L.STATUS_WORDS), dtype=torch.int32).pin_memory()
        _status_blocks.append(block)
        _status_pool.extend((row, row.numpy().view("uint32")) for row in block)
    while not _status_pool:
        oldest = next((p[0] for p in ([_pending.get(dev_idx)] + list(_pending.values())) if p), None)
        if oldest is None:  # every buffer is held by a captured graph: grow (outside any capture)
            block = torch.zeros((256, L.STATUS_WORDS), dtype=torch.int32).pin_memory()
            _status_blocks.append(block)
            _status_pool.extend((row, row.numpy().view("uint32")) for row in block)
            break
        _resolve(oldest, block=True)
    return _status_pool.pop()


class _AsyncFrame:
    """What stays behind of an asynchronous forward until its status words have been looked at.  Deliberately NOT the
    forward's state: that holds the frame's buffers (~200 MB at 1080p), and a host that runs many frames ahead of the
    device would keep all of them alive through the pending list — the allocator then grows by cudaMalloc, which
    synchronises (measured: 11 ms per step instead of 1.8)."""
    __slots__ = ("status", "stamp", "capacity", "dev_idx", "device", "num_rendered")

    def __init__(self, status, stamp, capacity, dev_idx, device):
        self.status, self.stamp, self.capacity, self.dev_idx, self.device = status, stamp, capacity, dev_idx, device
        self.num_rendered = None


def _resolve(frame, block: bool) -> bool:
    """Looks at the status words of an asynchronous forward (its _AsyncFrame, or the forward's state); True once they
    are known.  Raises on overflow."""
    state = None
    if isinstance(frame, dict):
        state, frame = frame, frame.get("async")
    if frame is None:
        return True
    st = frame.status
    if st is not None:
        tensor, words = st
        if int(words[L.STATUS_STAMP]) != frame.stamp:
            if not block:
                return False
            torch.cuda.synchronize(frame.device)  # the copy into pinned memory was enqueued behind the frame
            if int(words[L.STATUS_STAMP]) != frame.stamp:
                raise LsrError("asynchronous forward: status words did not arrive (was the frame captured in a CUDA "
                               "graph? use GraphedStep.check())")
        vals = [int(w) for w in words]
        frame.status = None
        _status_pool.append(st)
        pend = _pending.get(frame.dev_idx)
        if pend is not None:
            pend[:] = [x for x in pend if x is not frame]
        R, flags = vals[L.STATUS_NUM_RENDERED], vals[L.STATUS_FLAGS]
        frame.num_rendered = R
        _note_R(frame.dev_idx, R)
        if state is not None:
            state["num_rendered"] = R
        if flags & L.STATUS_FLAG_PREFILTERED:  # auxiliary.h:156-160 traps; here the frame is reported
            raise LsrError("Point is filtered although prefiltered is set. This shouldn't happen!")
        if flags & L.STATUS_FLAG_OVERFLOW:
            raise LsrError(f"asynchronous forward: the frame has {R} Gaussian/tile instances but its list was sized "
                           f"for {frame.capacity}; its images were filled with NaN on the device.  Repeat the step (the "
                           "next forward sizes its list from this count), raise capacity_margin, or call "
                           "reserve_instances().")
    if state is not None:
        state["num_rendered"] = frame.num_rendered
    return True


def check_async_status(device=None) -> None:
    """Waits for every outstanding asynchronous forward (of `device`, default all) and raises if one overflowed."""
    for idx in list(_pending):
        if device is not None:
            d = torch.device(device)
            if (d.index if d.index is not None else torch.cuda.current_device()) != idx:
                continue
        for state in list(_pending.get(idx, ())):
            _resolve(state, block=True)


def _poll_pending(dev_idx: int) -> None:
    for state in list(_pending.get(dev_idx, ())):
        if not _resolve(state, block=False):
            break  # stream order: later frames cannot be done either


def _present(t: Optional[torch.Tensor]) -> bool:
    return t is not None and t.numel() != 0


def _prep(t: Optional[torch.Tensor], device) -> Optional[torch.Tensor]:
    """`.contiguous()` as rasterize_points.cu:103-122 does (required: Camera.world_view_transform is a transposed
    view, scene/cameras.py:54); absent -> None."""
    if t is None:
        return None
    if t.dtype is torch.float32 and t.device == device and t.is_contiguous():  # the common case, kept cheap: this
        if t.numel() == 0:                                                     # runs a dozen times per frame
            return None
        # the kernels use 128-bit accesses: a contiguous view may start anywhere in its storage (x[1:]) -> re-pack
        return t if (t.data_ptr() & 15) == 0 else t.clone(memory_format=torch.contiguous_format)
    if not _present(t):
        return None
    if t.device != device or t.dtype != torch.float32:
        t = t.to(device=device, dtype=torch.float32)
    t = t.contiguous()
    return t if (t.data_ptr() & 15) == 0 else t.clone(memory_format=torch.contiguous_format)


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _make_frame(rs: GaussianRasterizationSettings, P: int, M: int, bg, view, proj, campos) -> L.Frame:
    f = L.Frame()
    f.P, f.M, f.D = P, M, int(rs.sh_degree)
    f.width, f.height = int(rs.image_width), int(rs.image_height)
    f.tanfovx, f.tanfovy, f.scale_modifier = float(rs.tanfovx), float(rs.tanfovy), float(rs.scale_modifier)
    f.prefiltered, f.debug, f.include_feature = int(bool(rs.prefiltered)), int(bool(rs.debug)), int(
        bool(rs.include_feature))
    f.bg, f.viewmatrix, f.projmatrix, f.campos = bg.data_ptr(), view.data_ptr(), proj.data_ptr(), campos.data_ptr()
    return f


def _make_inputs(means3D, sh, colors, lang, opac, scales, rots, cov3D) -> L.Inputs:
    i = L.Inputs()
    for name, t in (("means3D", means3D), ("shs", sh), ("colors_precomp", colors), ("language_feature", lang),
                    ("opacities", opac), ("scales", scales), ("rotations", rots), ("cov3D_precomp", cov3D)):
        setattr(i, name, None if t is None else t.data_ptr())
    return i


def _get_scratch(device, stream_ptr: int, nbytes: int) -> torch.Tensor:
    key = (device.index if device.index is not None else torch.cuda.current_device(), stream_ptr)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(int(nbytes * 1.25) + 1024, dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf


_size_cache: Dict[tuple, int] = {}


def _sized(fn, *args) -> int:
    """Memoised size queries (lsr_geom_bytes & co. are pure functions of their arguments)."""
    key = (fn.__name__,) + args
    v = _size_cache.get(key)
    if v is None:
        if len(_size_cache) > 256:
            _size_cache.clear()
        v = _size_cache[key] = int(fn(*args))
    return v


class _NoContext:
    def __enter__(self):
        return None

    def __exit__(self, *a):
        return False


_NO_CONTEXT = _NoContext()


def _on_device(device):
    """`with torch.cuda.device(device)` only when the current device differs (the guard costs ~10 us per use)."""
    idx = device.index
    if idx is None or idx == torch.cuda.current_device():
        return _NO_CONTEXT
    return torch.cuda.device(device)


def _check(rc: int, what: str):
    if rc != L.LSR_OK:
        raise LsrError(f"{what}: {L.last_error()} (status {rc})")


def _forward_impl(rs: GaussianRasterizationSettings, means3D, sh, colors, lang, opac, scales, rots, cov3D,
                  keep_scratch: bool = False):
    """Runs lsr_forward.  Returns (color, feature, radii, state) where state holds what backward needs."""
    lib = L.load()
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")  # rasterize_points.cu:59-61
    if not means3D.is_cuda:
        raise LsrError("langsplat_b200 has no CPU path: means3D must be a CUDA tensor")
    device = means3D.device
    P = means3D.size(0)
    H, W = int(rs.image_height), int(rs.image_width)
    means3D = _prep(means3D, device)
    sh, colors, opac = _prep(sh, device), _prep(colors, device), _prep(opac, device)
    scales, rots, cov3D = _prep(scales, device), _prep(rots, device), _prep(cov3D, device)
    include_feature = bool(rs.include_feature)
    lang = _prep(lang, device) if include_feature else None
    bg, view, proj, campos = (_prep(rs.bg, device), _prep(rs.viewmatrix, device), _prep(rs.projmatrix, device),
                              _prep(rs.campos, device))
    M = sh.size(1) if sh is not None else 0
    frame = _make_frame(rs, P, M, bg, view, proj, campos)
    inputs = _make_inputs(means3D, sh, colors, lang, opac, scales, rots, cov3D)

    color = torch.empty((3, H, W), dtype=torch.float32, device=device)
    feature = (torch.empty((3, H, W), dtype=torch.float32, device=device) if include_feature else torch.zeros(
        (1,), dtype=torch.float32, device=device))
    radii = torch.empty((P,), dtype=torch.int32, device=device)
    geom = torch.empty((_sized(lib.lsr_geom_bytes, P),), dtype=torch.uint8, device=device)
    image = torch.empty((_sized(lib.lsr_image_bytes, W, H),), dtype=torch.uint8, device=device)
    stream = torch.cuda.current_stream(device).cuda_stream
    dev_idx = device.index if device.index is not None else torch.cuda.current_device()
    global _capacity_retries
    recent = _recent_R.get(dev_idx)
    if _AsyncConfig.enabled and recent and not rs.debug:
        _poll_pending(dev_idx)  # (raises if an earlier frame on this device outgrew its list)
        capacity = max(int(max(recent) * _AsyncConfig.capacity_margin), _INITIAL_CAPACITY_FLOOR)
        capacity = (capacity + 0xFFFF) & ~0xFFFF
        status = _take_status(dev_idx)
        status[1][L.STATUS_STAMP] = 0
        _stamp[0] = _stamp[0] % 0xFFFFFFF0 + 1
        with _on_device(device):
            point_list = torch.empty((_sized(lib.lsr_list_bytes, capacity),), dtype=torch.uint8, device=device)
            need = _sized(lib.lsr_scratch_bytes, P, W, H, capacity)
            scratch = (torch.empty((need,), dtype=torch.uint8, device=device) if keep_scratch else _get_scratch(
                device, stream, need))
            rc = lib.lsr_forward_nosync(ctypes.byref(frame), ctypes.byref(inputs), _ptr(color), _ptr(feature),
                                        _ptr(radii), _ptr(geom), _ptr(image), _ptr(point_list), capacity, _ptr(scratch),
                                        scratch.numel(), _stamp[0], ctypes.c_void_p(status[0].data_ptr()),
                                        ctypes.c_void_p(stream))
        if rc != L.LSR_OK:
            _status_pool.append(status)
        _check(rc, "lsr_forward_nosync")
        pending = _AsyncFrame(status, _stamp[0], capacity, dev_idx, device)
        state = dict(frame=frame, inputs=inputs, geom=geom, image=image, point_list=point_list, num_rendered=None,
                     capacity=capacity, keep=(means3D, sh, colors, lang, opac, scales, rots, cov3D, bg, view, proj,
                                              campos),
                     radii=radii, P=P, M=M, W=W, H=H, device=device, dev_idx=dev_idx, stream=stream)
        state["async"] = pending
        _pending.setdefault(dev_idx, []).append(pending)
        if keep_scratch:
            state["scratch"] = scratch
        return color, feature, radii, state
    capacity = max(int(_last_R.get(dev_idx, 0) * 1.25), _CAPACITY_PER_GAUSSIAN * P, _INITIAL_CAPACITY_FLOOR)
    capacity = (capacity + 0xFFFF) & ~0xFFFF  # 64 Ki steps: stable buffer sizes from frame to frame
    nr = ctypes.c_int64(0)
    with _on_device(device):
        while True:
            point_list = torch.empty((_sized(lib.lsr_list_bytes, capacity),), dtype=torch.uint8, device=device)
            need = _sized(lib.lsr_scratch_bytes, P, W, H, capacity)
            scratch = (torch.empty((need,), dtype=torch.uint8, device=device) if keep_scratch else _get_scratch(
                device, stream, need))
            rc = lib.lsr_forward(ctypes.byref(frame), ctypes.byref(inputs), _ptr(color), _ptr(feature), _ptr(radii),
                                 _ptr(geom), _ptr(image), _ptr(point_list), capacity, _ptr(scratch), scratch.numel(),
                                 ctypes.byref(nr), ctypes.c_void_p(stream))
            if rc == L.LSR_ERR_CAPACITY:
                _capacity_retries += 1
                capacity = int(nr.value * 1.1) + 1024
                continue
            break
    _check(rc, "lsr_forward")
    _note_R(dev_idx, int(nr.value))
    state = dict(frame=frame, inputs=inputs, geom=geom, image=image, point_list=point_list, num_rendered=int(nr.value),
                 capacity=capacity, keep=(means3D, sh, colors, lang, opac, scales, rots, cov3D, bg, view, proj, campos),
                 radii=radii, P=P, M=M, W=W, H=H, device=device, dev_idx=dev_idx, stream=stream)
    if keep_scratch:
        state["scratch"] = scratch
    return color, feature, radii, state


# When a list is installed here (langsplat_b200.dist.factored_sh_gradients), a backward that owes an SH gradient does
# not materialise dL_dsh [P,16,3]; it appends that view's factor — (clamp-masked dL_dcolor [P,3], campos [3],
# sh_degree) — and returns no gradient for `shs`.  dL_dsh of a view is the rank-1 product basis(dir) (x) factor, so a
# multi-view / data-parallel step can exchange and accumulate 12-byte factors instead of 192-byte products and form
# the sum once (lsr_sh_gradient_from_views).
_sh_factor_sink = None


class ShFactorList(list):
    """Default sink: one fresh [P,3] tensor per view; entries are (factor, campos, sh_degree).  Any object with this
    `new_factor` method can serve as a sink (langsplat_b200.dist.PeerFactorExchange hands out peer-visible memory)."""

    def new_factor(self, P: int, device, campos: torch.Tensor, degree: int) -> torch.Tensor:
        t = torch.empty((P, 3), dtype=torch.float32, device=device)
        self.append((t, campos, degree))
        return t


def _grad_layout(P: int, n_sh: int, run_pre: bool, use_acc: bool, need_colors: bool, external_params: bool):
    """Offsets (in floats) of the backward's buffers inside ONE allocation: {name: (in the caller's parameter block?,
    offset, length)}, the floats to allocate, the floats of the parameter block.
    The parameter gradients come first and back to back (rotations, sh, means3D, scales, language feature, opacity — the
    two that the kernels access as float4 lead, so they stay 16-byte aligned for any P), so that a data-parallel
    caller can all-reduce them as ONE flat message (langsplat_b200.dist.coalesce); with `external_params` that block
    lives in memory the caller provides.  The internal buffers follow from the next 64-byte boundary, widest access
    first: the accumulation record (64-byte vector reductions), dL_dconic (float4), dL_dcov3D (float2), then the
    triplets.  tests/test_layout.py checks the alignment of every vector access for odd P."""
    sizes = [("rots", 4 if run_pre else 0), ("sh", n_sh), ("means3D", 3 if run_pre else 0),
             ("scales", 3 if run_pre else 0), ("feat", 3), ("opac", 1), ("|", 0), ("acc", 16 if use_acc else 0),
             ("conic", 0 if use_acc else 4), ("cov3D", 6 if run_pre else 0), ("means2D", 3),
             ("colors", 3 if need_colors else 0)]
    places, off, in_lead = {}, 0, external_params
    n_params = sum(n for _, n in sizes[:sizes.index(("|", 0))]) * P
    for name, n in sizes:
        if name == "|":
            off = 0 if external_params else (off + 15) // 16 * 16  # (torch allocations start 512-byte aligned)
            in_lead = False
            continue
        places[name] = (in_lead, off, n * P)
        off += n * P
    return places, off + 16, n_params


def _backward_impl(state, grad_color, grad_feature, need_mask: int, want_sh: bool, sh_factor_out=None):
    lib = L.load()
    _resolve(state, block=False)  # asynchronous forward: raises if its instance list is known to have overflowed
    P, M, W, H, device = state["P"], state["M"], state["W"], state["H"], state["device"]
    (means3D, sh, colors, lang, opac, scales, rots, cov3D, *_rest) = state["keep"]
    f32 = dict(dtype=torch.float32, device=device)
    include_feature = bool(state["frame"].include_feature)
    gc = grad_color.contiguous() if grad_color is not None else None
    gf = grad_feature.contiguous() if (grad_feature is not None and include_feature) else None
    geom_need = bool(need_mask & L.NEED_GEOMETRY)
    run_pre = geom_need or (sh is not None and want_sh and gc is not None)
    # One allocation for everything (_grad_layout).  The library zero-fills what the blend kernel accumulates into
    # (adjacent ranges merge into one memset).
    factored = sh_factor_out is not None and run_pre and sh is not None and gc is not None
    n_sh = M * 3 if (run_pre and sh is not None and not factored) else 0
    # The blend kernel accumulates into ONE packed 64-byte record per Gaussian ("acc", lsr_grads::blend_acc: 16-byte
    # vector reductions instead of twelve scalar atomics into five tensors); the kernel behind it writes the tensors.
    # No dL_dconic buffer then, and no dL_dcolors unless the caller gets it back (colors_precomp).  LSR_BLEND_ACC=0
    # keeps the five accumulators (A/B, and what the one-phase kernel of LSR_BWD_V1=1 needs).
    # Only when the per-Gaussian kernel follows (it reads the record for free): a feature-only backward would pay an
    # extra unpack pass for nothing (GPU call r2s, C3: blend kernel 484 us either way, + 10 us of unpack and zero-fill).
    use_acc = (_USE_BLEND_ACC == 2 or (_USE_BLEND_ACC == 1 and run_pre)) and bool(lib.lsr_blend_acc_supported())
    need_colors = not use_acc or sh is None
    # A sink may provide the memory of the parameter gradients (PeerFactorExchange: peer-visible staging memory, so
    # that the all-reduce reads them in place); the internal accumulators then come from their own allocation.
    n_params = _grad_layout(P, n_sh, run_pre, use_acc, need_colors, False)[2]
    lead = sh_factor_out.new_flat(n_params, device) if factored and hasattr(sh_factor_out, "new_flat") else None
    places, total, _ = _grad_layout(P, n_sh, run_pre, use_acc, need_colors, lead is not None)
    flat = torch.empty((total,), **f32)
    part = {}
    for name, (in_lead, off, n) in places.items():
        part[name] = (lead if in_lead else flat)[off:off + n] if n else None
    d_conic = part["conic"].view(P, 4) if part["conic"] is not None else None
    d_means2D = part["means2D"].view(P, 3)
    d_opac = part["opac"].view(P, 1)
    d_colors = part["colors"].view(P, 3) if part["colors"] is not None else None
    d_feat = part["feat"].view(P, 3)
    g = L.Grads()
    g.dL_dmeans2D, g.dL_dopacity, g.dL_dlanguage_feature = d_means2D.data_ptr(), d_opac.data_ptr(), d_feat.data_ptr()
    g.dL_dconic = d_conic.data_ptr() if d_conic is not None else None
    g.dL_dcolors = d_colors.data_ptr() if d_colors is not None else None
    if use_acc:
        g.blend_acc = part["acc"].data_ptr()
        assert g.blend_acc % 64 == 0
    d_means3D = d_cov3D = d_sh = d_scales = d_rots = None
    if run_pre:
        d_means3D = part["means3D"].view(P, 3)
        d_cov3D = part["cov3D"].view(P, 6)
        g.dL_dmeans3D, g.dL_dcov3D = d_means3D.data_ptr(), d_cov3D.data_ptr()
        if sh is not None and not factored:
            d_sh = part["sh"].view(P, M, 3)
            g.dL_dsh = d_sh.data_ptr()
        if factored:
            masked = sh_factor_out.new_factor(P, device, state["keep"][11], int(state["frame"].D))
            g.dL_dcolors_masked = masked.data_ptr()
        d_scales = part["scales"].view(P, 3)
        d_rots = part["rots"].view(P, 4)
        g.dL_dscales, g.dL_drotations = d_scales.data_ptr(), d_rots.data_ptr()
    stream = torch.cuda.current_stream(device).cuda_stream
    with _on_device(device):
        rc = lib.lsr_backward(ctypes.byref(state["frame"]), ctypes.byref(state["inputs"]), _ptr(state["radii"]),
                              _ptr(state["geom"]), _ptr(state["image"]), _ptr(state["point_list"]), _ptr(gc), _ptr(gf),
                              need_mask | (L.NEED_GEOMETRY if run_pre else 0), ctypes.byref(g),
                              ctypes.c_void_p(stream))
    _check(rc, "lsr_backward")
    return dict(means3D=d_means3D, means2D=d_means2D, sh=d_sh, colors=d_colors, language_feature=d_feat,
                opacities=d_opac, scales=d_scales, rotations=d_rots, cov3D=d_cov3D, conic=d_conic)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, language_feature_precomp, opacities, scales, rotations,
                cov3Ds_precomp, raster_settings):
        args = (means3D, sh, colors_precomp, language_feature_precomp, opacities, scales, rotations, cov3Ds_precomp)
        if raster_settings.debug:  # RAST/diff_gaussian_rasterization/__init__.py:87-95
            cpu_args = cpu_deep_copy_tuple((raster_settings.bg,) + args + tuple(raster_settings))
            try:
                color, feature, radii, state = _forward_impl(raster_settings, *args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            color, feature, radii, state = _forward_impl(raster_settings, *args)
        ctx.raster_settings = raster_settings
        ctx.num_rendered = state["num_rendered"]  # None until known when the forward ran asynchronously
        ctx.state = state
        # The kernels' view of the inputs travels in `state` (device pointers of the contiguous tensors); the tensors
        # themselves go through save_for_backward like the reference's (RAST/diff_gaussian_rasterization/
        # __init__.py:102), so that autograd's version counters catch an in-place update between forward and backward
        # (the backward combines the forward's geometry records with these inputs) and saved-tensor hooks see them.
        ctx.save_for_backward(*args)
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(radii)
        return color, feature, radii

    @staticmethod
    def backward(ctx, grad_out_color, grad_out_language_feature, _):
        state = ctx.state
        ctx.saved_tensors  # raises if an input was modified in place since the forward (version counters)
        need = ctx.needs_input_grad  # (means3D, means2D, sh, colors, lang, opac, scales, rots, cov3D, settings)
        mask = 0
        if need[0] or need[5] or need[6] or need[7] or need[8]:
            mask |= L.NEED_GEOMETRY
        if need[2] or need[3]:
            mask |= L.NEED_COLOR
        if need[4]:
            mask |= L.NEED_FEATURE
        if need[1]:
            mask |= L.NEED_MEANS2D
        include_feature = bool(ctx.raster_settings.include_feature)
        if grad_out_color is None and (grad_out_language_feature is None or not include_feature):
            return (None,) * 10
        want_sh = bool(need[2] or need[0])
        if ctx.raster_settings.debug:  # :139-147
            try:
                g = _backward_impl(state, grad_out_color, grad_out_language_feature, mask, want_sh)
            except Exception as ex:
                torch.save(cpu_deep_copy_tuple((grad_out_color, grad_out_language_feature)), "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            g = _backward_impl(state, grad_out_color, grad_out_language_feature, mask, want_sh,
                               _sh_factor_sink if need[2] else None)
        lang_grad = None
        if need[4]:
            lang_grad = g["language_feature"] if include_feature else torch.zeros(
                (1,), dtype=torch.float32, device=state["device"])  # rasterize_points.cu:176-178
        return (
            g["means3D"] if need[0] else None,
            g["means2D"] if need[1] else None,
            g["sh"] if need[2] else None,
            g["colors"] if need[3] else None,
            lang_grad,
            g["opacities"] if need[5] else None,
            g["scales"] if need[6] else None,
            g["rotations"] if need[7] else None,
            g["cov3D"] if need[8] else None,
            None,
        )


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, language_feature_precomp, opacities, scales, rotations,
                        cov3Ds_precomp, raster_settings):  # RAST/diff_gaussian_rasterization/__init__.py:21-44
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, language_feature_precomp, opacities, scales,
                                     rotations, cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):  # RAST/diff_gaussian_rasterization/__init__.py:181-233
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """Boolean mask of points passing the frustum test (p_view.z > 0.2), :186-195."""
        with torch.no_grad():
            rs = self.raster_settings
            lib = L.load()
            if not positions.is_cuda:
                raise LsrError("langsplat_b200 has no CPU path: positions must be a CUDA tensor")
            device = positions.device
            pos = _prep(positions, device)
            P = positions.size(0)
            present = torch.zeros((P,), dtype=torch.bool, device=device)
            if P:
                view, proj = _prep(rs.viewmatrix, device), _prep(rs.projmatrix, device)
                with torch.cuda.device(device):
                    rc = lib.lsr_mark_visible(P, _ptr(pos), _ptr(view), _ptr(proj), _ptr(present),
                                              ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream))
                _check(rc, "lsr_mark_visible")
        return present

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, language_feature_precomp=None,
                scales=None, rotations=None, cov3D_precomp=None):
        raster_settings = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception(_MSG_COLORS)
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                (scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception(_MSG_COV)
        empty = torch.Tensor([])
        shs = empty if shs is None else shs
        colors_precomp = empty if colors_precomp is None else colors_precomp
        language_feature_precomp = empty if language_feature_precomp is None else language_feature_precomp
        scales = empty if scales is None else scales
        rotations = empty if rotations is None else rotations
        cov3D_precomp = empty if cov3D_precomp is None else cov3D_precomp
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, language_feature_precomp, opacities, scales,
                                   rotations, cov3D_precomp, raster_settings)


    def forward_feature_levels(self, means3D, opacities, language_features, shs=None, colors_precomp=None,
                               scales=None, rotations=None, cov3D_precomp=None):
        """Render several language-feature levels of the SAME geometry in one go (inference only, no autograd):
        `language_features` is a sequence of [P,3] tensors.  LangSplat trains its three SAM levels on frozen geometry,
        and render.py / process.sh render them back to back (render.py:35-52, process.sh:19-32); here the preprocess
        and the binning run once and every further level costs one feature-only blend (lsr_forward_feature_level).
        Returns (color, [feature image per level], radii); each feature image equals what a separate `forward` call
        with that level would return.  Extension over the reference surface."""
        rs = self.raster_settings
        if not rs.include_feature:
            raise LsrError("forward_feature_levels needs raster_settings.include_feature=True")
        levels = list(language_features)
        if not levels:
            raise LsrError("forward_feature_levels needs at least one language feature tensor")
        if (shs is None) == (colors_precomp is None):
            raise Exception(_MSG_COLORS)
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                (scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception(_MSG_COV)
        lib = L.load()
        with torch.no_grad():
            color, feat0, radii, state = _forward_impl(rs, means3D, shs, colors_precomp, levels[0], opacities, scales,
                                                       rotations, cov3D_precomp)
            feats = [feat0]
            device, P, W, H = state["device"], state["P"], state["W"], state["H"]
            stream = ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
            for lang in levels[1:]:
                lang = _prep(lang, device)
                if lang is None or lang.numel() != P * 3:
                    raise LsrError("every language feature level must have shape (num_points, 3)")
                out = torch.empty((3, H, W), dtype=torch.float32, device=device)
                with torch.cuda.device(device):
                    rc = lib.lsr_forward_feature_level(P, W, H, _ptr(radii), _ptr(state["geom"]), _ptr(state["image"]),
                                                       _ptr(state["point_list"]), _ptr(lang), _ptr(out), stream)
                _check(rc, "lsr_forward_feature_level")
                feats.append(out)
        return color, feats, radii


# ---- test / bench helpers (not part of the reference surface) ----
def debug_arrays(state) -> Dict[str, torch.Tensor]:
    """Typed views of the arrays inside a forward's geom / image / scratch buffers (include/lsr_api.h LSR_ARR_*)."""
    lib = L.load()
    _resolve(state, block=True)  # an asynchronous forward learns its instance count here
    P, W, H, cap, R = state["P"], state["W"], state["H"], state["capacity"], state["num_rendered"]
    T = ((W + 15) // 16) * ((H + 15) // 16)

    def view(buf, which, dtype, count):
        off = lib.lsr_array_offset(which, P, W, H, cap)
        nbytes = count * torch.empty((), dtype=dtype).element_size()
        return buf[off:off + nbytes].view(dtype)

    geom, image = state["geom"], state["image"]
    out = dict(
        rec=view(geom, L.ARR_GEOM_REC, torch.float32, P * 12).view(P, 12),
        depths=view(geom, L.ARR_GEOM_DEPTH, torch.float32, P),
        tiles_touched=view(geom, L.ARR_GEOM_TILES, torch.int32, P),
        point_offsets=view(geom, L.ARR_GEOM_OFFSETS, torch.int32, P),
        cov3D=view(geom, L.ARR_GEOM_COV3D, torch.float32, P * 6).view(P, 6),
        clamped=view(geom, L.ARR_GEOM_CLAMPED, torch.uint8, P),
        final_T=view(image, L.ARR_IMG_FINAL_T, torch.float32, W * H).view(H, W),
        n_contrib=view(image, L.ARR_IMG_NCONTRIB, torch.int32, W * H).view(H, W),
        ranges=view(image, L.ARR_IMG_RANGES, torch.int32, T * 2).view(T, 2),
        point_list=state["point_list"][:R * 4].view(torch.int32),
    )
    if "scratch" in state:
        sc = state["scratch"]
        # (the unsorted keys are overwritten by the sort's ping-pong; the sorted keys are the same multiset)
        keys = torch.empty((R,), dtype=torch.int64, device=sc.device)
        out["depth_order"] = view(sc, L.ARR_SCR_DEPTH_ORDER, torch.int32, P)
        rc = lib.lsr_debug_sorted_keys(_ptr(sc), _ptr(geom), P, cap, R, _ptr(keys),
                                       ctypes.c_void_p(torch.cuda.current_stream(sc.device).cuda_stream))
        _check(rc, "lsr_debug_sorted_keys")
        out["keys"] = keys
    return out
