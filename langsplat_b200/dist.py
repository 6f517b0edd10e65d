"""View-sharded data parallelism (BASELINE.json config 4).

The reference is single-GPU (one camera per step, train.py:84-92).  The natural shard of the rasterizer is the camera
view: Gaussians are replicated, rank r of n renders views {r, r+n, ...} of the step and runs its own backward; the one
exchange step is a sum-allreduce of the Gaussian-parameter gradients (NCCL over NVLink/NVSwitch on the GPU box, gloo
in the CPU tests).  No collective touches the rasterizer's data path itself.
"""
from __future__ import annotations

import contextlib
import ctypes
import os
import sys
from typing import Dict, Iterable, List, Optional, Sequence

import torch
import torch.distributed as dist


def views_for_rank(n_views: int, rank: int, world_size: int) -> List[int]:
    """Round-robin assignment of the step's camera views to ranks."""
    return list(range(rank, n_views, world_size))


def allreduce_gradients(grads: Sequence[Optional[torch.Tensor]], group=None, async_op: bool = False):
    """Sum-allreduce every non-None gradient tensor in place.  One collective per tensor, issued back to back so NCCL
    pipelines them; the per-attribute order (xyz, features, opacity, scaling, rotation, language feature) is the order
    the per-Gaussian backward finishes them in.  Returns the list of work handles when async_op."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return []
    handles = []
    for g in coalesce(grads):
        h = dist.all_reduce(g, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
        if async_op:
            handles.append(h)
    return handles


def coalesce(grads: Sequence[Optional[torch.Tensor]]) -> List[torch.Tensor]:
    """Merges gradient tensors that sit back to back in one allocation into a single flat view, so that they travel
    as one collective.  The rasterizer's backward returns the parameter gradients (means3D, sh, opacity, scales,
    rotations, language feature) as adjacent slices of one buffer precisely for this."""
    live = [g for g in grads if g is not None]
    out: List[torch.Tensor] = []
    by_store = {}
    for g in live:
        if g.is_contiguous() and g.dtype == torch.float32:
            by_store.setdefault(g.untyped_storage().data_ptr(), []).append(g)
        else:
            out.append(g)
    for gs in by_store.values():
        gs.sort(key=lambda t: t.data_ptr())
        run = [gs[0]]
        for t in gs[1:]:
            prev = run[-1]
            if t.data_ptr() == prev.data_ptr() + prev.numel() * 4:
                run.append(t)
            else:
                out.append(_flat(run))
                run = [t]
        out.append(_flat(run))
    return out


def _flat(run: List[torch.Tensor]) -> torch.Tensor:
    if len(run) == 1:
        return run[0]
    first = run[0]
    n = sum(t.numel() for t in run)
    return first.new_empty(0).set_(first.untyped_storage(), first.storage_offset(), (n,), (1,))


def accumulate_view_gradients(params: Dict[str, torch.Tensor], render_and_loss, views: Iterable[int]) -> None:
    """Runs `render_and_loss(view_index) -> scalar loss` for each local view and accumulates .grad on `params`
    (plain autograd accumulation; gradients of different views add, as they would in a batch)."""
    for v in views:
        loss = render_and_loss(v)
        loss.backward()


def reduce_view_statistics(stat_sum: torch.Tensor, stat_count: torch.Tensor, max_radii: torch.Tensor, group=None):
    """Densification statistics must be reduced as statistics, not derived from summed gradients
    (scene/gaussian_model.py:480-482, train.py:122-133): sum of per-view ||grad means2D||, visibility counts, max radii."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    dist.all_reduce(stat_sum, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(stat_count, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(max_radii, op=dist.ReduceOp.MAX, group=group)


# ---- factored SH gradients --------------------------------------------------------------------------------------
# 81 % of the parameter-gradient bytes of an RGB step are dL_dsh (192 of 236 B per Gaussian), yet one view's dL_dsh is
# rank-1 per Gaussian: basis(normalize(mean - campos)) (x) g with g = the clamp-masked dL_dcolor (RAST backward.cu:
# 31-129).  Inside `factored_sh_gradients()` the rasterizer's backward hands out the 12-byte factor g instead of the
# product; `finish_sh_gradients` all-gathers the factors of every rank's views (12 B per Gaussian and view instead of
# an all-reduce of 192 B) and forms sum_v basis(dir_v) (x) g_v once with lsr_sh_gradient_from_views.  The result equals
# the all-reduced dL_dsh up to float summation order.  On one GPU it still pays for multi-view steps: no per-view
# [P,16,3] gradient is written and re-accumulated.

@contextlib.contextmanager
def factored_sh_gradients(sink=None):
    """Context in which GaussianRasterizer backward passes hand out per-view SH-gradient factors instead of dL_dsh.
    Yields the sink: by default a list of (factor [P,3], campos, sh_degree), one entry per backward; pass
    `PeerFactorExchange.begin_step()` to have the factors written straight into peer-visible memory."""
    from . import rasterizer as R
    if sink is None:
        sink = R.ShFactorList()
    prev = R._sh_factor_sink
    R._sh_factor_sink = sink
    try:
        yield sink
    finally:
        R._sh_factor_sink = prev


def gather_sh_factors(factors, P: int, device, group=None):
    """(g [V_all, P, 3], campos [V_all, 3], sh_degree) over the views of all ranks, rank-major (all_gather order).
    Every rank must contribute the same number of views.  Two collectives: 12 B per Gaussian and view, and the
    camera positions."""
    g_local = (torch.stack([f[0] for f in factors]) if factors
               else torch.zeros((0, P, 3), dtype=torch.float32, device=device))
    c_local = (torch.stack([f[1].reshape(3).to(device, torch.float32) for f in factors]) if factors
               else torch.zeros((0, 3), dtype=torch.float32, device=device))
    degree = int(factors[0][2]) if factors else 0
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1):
        return g_local.contiguous(), c_local.contiguous(), degree
    world = dist.get_world_size(group)
    g_all = torch.empty((world * g_local.shape[0], P, 3), dtype=torch.float32, device=device)
    c_all = torch.empty((world * c_local.shape[0], 3), dtype=torch.float32, device=device)
    dist.all_gather_into_tensor(g_all, g_local.contiguous(), group=group)
    dist.all_gather_into_tensor(c_all, c_local.contiguous(), group=group)
    return g_all, c_all, degree


def finish_sh_gradients(factors, means3D: torch.Tensor, n_coeffs: int, group=None) -> Optional[torch.Tensor]:
    """dL_dsh [P, n_coeffs, 3] summed over the views of ALL ranks from the factors collected on this rank
    (lsr_sh_gradient_from_views).  Returns None if there is no process group and no view produced a factor."""
    from . import _lib as L
    from .rasterizer import LsrError
    distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    if not factors and not distributed:
        return None
    P = means3D.shape[0]
    dev = means3D.device
    if not means3D.is_cuda:
        raise LsrError("finish_sh_gradients: langsplat_b200 has no CPU path")
    g_all, c_all, degree = gather_sh_factors(factors, P, dev, group)
    lib = L.load()
    out = torch.empty((P, n_coeffs, 3), dtype=torch.float32, device=dev)
    m3 = means3D.detach().contiguous()
    with torch.cuda.device(dev):
        rc = lib.lsr_sh_gradient_from_views(P, n_coeffs, degree, g_all.shape[0], ctypes.c_void_p(m3.data_ptr()),
                                            ctypes.c_void_p(c_all.data_ptr()), ctypes.c_void_p(g_all.data_ptr()),
                                            ctypes.c_void_p(out.data_ptr()),
                                            ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    if rc != L.LSR_OK:
        raise LsrError(L.last_error())
    return out


# ---- the exchange fused into the kernel: factors read in place over NVLink ------------------------------------------
class _DevicePtr:
    """Lets torch wrap device memory it did not allocate (the peer buffer) as a tensor."""

    def __init__(self, ptr: int, n_floats: int):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (ptr, False), "version": 2}


class PeerFactorExchange:
    """NVLink exchange of the SH-gradient factors without a collective: every rank's backward writes its views'
    factors into a buffer the other ranks have mapped (lsr_peer_alloc / lsr_peer_open, CUDA IPC); after one
    stream-ordered barrier kernel (lsr_peer_barrier) the SH-gradient kernel reads all world x views_per_rank factors in
    place — its loads of the other ranks' factors ARE the transfer, overlapped tile by tile with the accumulation and
    the dL_dsh stores (lsr_sh_gradient_from_view_ptrs).  No all-gather, no staging copy, and NCCL stays free to
    all-reduce the remaining gradients at the same time.

    Optionally (grad_floats_per_gaussian > 0, `exchange`) the remaining gradients are summed over peer memory as well,
    by a two-shot all-reduce kernel.  Measured at N=2, 56 MB: that kernel alone 85 us against NCCL's 127 us, but it
    and the SH-gradient kernel contend for the same links and SMs and end up back to back (185 us for the whole
    exchange), whereas NCCL's copy-engine-like traffic overlaps the SH-gradient kernel (136 us) — so `finish` + NCCL
    is the default and `exchange` the alternative (profiles/tools/exchange_times.py, bench.py --peer-allreduce).

    Two slots alternate by step parity: a rank can only get two steps ahead of a peer by passing the barrier in
    between, which that peer enqueues after the kernel that read the slot, so no slot is overwritten while in use.
    Every rank must run the same number of views per step, on the same stream from step to step."""

    def __init__(self, P: int, views_per_rank: int, device, group=None, grad_floats_per_gaussian: int = 0,
                 device_epochs: bool = False):
        """device_epochs: the barriers count their epochs on the device (lsr_peer_sync_auto) instead of taking them
        from this object, so that a step with its exchange can be captured in a CUDA graph and replayed; the payload
        slot of a step is then chosen by the caller (`begin_step(slot=...)`: capture one graph per slot and replay
        them alternately — a rank may be a step ahead of a peer that still reads the previous step's factors)."""
        from . import _lib as L
        from .rasterizer import LsrError
        self._L, self._lib, self._err = L, L.load(), LsrError
        self.P, self.V, self.device, self.group = int(P), int(views_per_rank), torch.device(device), group
        on = dist.is_available() and dist.is_initialized()
        self.rank = dist.get_rank(group) if on else 0
        self.world = dist.get_world_size(group) if on else 1
        if self.world > L.MAX_PEERS:
            raise LsrError(f"PeerFactorExchange: at most {L.MAX_PEERS} ranks")
        # one slot, in floats: the views' factors | 4 floats of camera position per view | (64-float aligned) this
        # rank's other gradients, input of the two-shot all-reduce | its output, written by all ranks
        self.cam_off = self.V * self.P * 3
        self.grad_cap = (int(grad_floats_per_gaussian) * self.P + 63) // 64 * 64
        self.in_off = (self.cam_off + 4 * self.V + 63) // 64 * 64
        self.out_off = self.in_off + self.grad_cap
        slot_floats = self.out_off + self.grad_cap
        self.slot_bytes = slot_floats * 4
        total = L.PEER_HEADER_BYTES + 2 * self.slot_bytes
        own, handle = ctypes.c_void_p(), (ctypes.c_uint8 * L.PEER_HANDLE_BYTES)()
        with torch.cuda.device(self.device):
            rc = self._lib.lsr_peer_alloc(total, ctypes.byref(own), handle)
        self.own = own.value if rc == L.LSR_OK else None
        self.bases: List[Optional[int]] = [None] * self.world
        self.bases[self.rank] = self.own
        self._opened: List[int] = []
        if self.world > 1:  # a rank whose allocation failed still takes part, so that nobody waits for it
            handles: List = [None] * self.world
            dist.all_gather_object(handles, bytes(handle) if rc == L.LSR_OK else None, group=group)
            if any(h is None for h in handles):
                if self.own is not None:
                    self._lib.lsr_peer_free(ctypes.c_void_p(self.own))
                    self.own = None
                raise LsrError("PeerFactorExchange: lsr_peer_alloc failed on a rank: " + L.last_error())
            with torch.cuda.device(self.device):
                for r in range(self.world):
                    if r != self.rank:
                        h = (ctypes.c_uint8 * L.PEER_HANDLE_BYTES).from_buffer_copy(handles[r])
                        p = ctypes.c_void_p()
                        self._check(self._lib.lsr_peer_open(h, ctypes.byref(p)))
                        self.bases[r] = p.value
                        self._opened.append(p.value)
        self._check(rc)
        self._payload = torch.as_tensor(_DevicePtr(self.own + L.PEER_HEADER_BYTES, 2 * slot_floats),
                                        device=self.device).view(2, slot_floats)
        self._headers = (ctypes.c_void_p * self.world)(*self.bases)
        self.epoch, self._count, self._degree, self._barriers = 0, 0, 0, 0
        self._slot, self.device_epochs = 0, bool(device_epochs)
        self._side = torch.cuda.Stream(device=self.device)
        # the sticky timed-out word of the own header, mirrored into pinned host memory behind every exchange and
        # looked at (without synchronising) when the next step begins
        self._flag_dev = torch.as_tensor(_DevicePtr(self.own, L.PEER_HEADER_BYTES // 4), device=self.device)[
            L.MAX_PEERS:L.MAX_PEERS + 1]
        self._flag_host = torch.zeros(1, dtype=torch.float32).pin_memory()
        ms = os.environ.get("LSR_PEER_TIMEOUT_MS")
        if ms:
            self._check(self._lib.lsr_peer_set_timeout_ms(int(ms)))

    def _check(self, rc: int) -> None:
        if rc != self._L.LSR_OK:
            raise self._err(self._L.last_error())

    # -- sink protocol (rasterizer._backward_impl) --
    def _sync(self, mode: int, stream) -> None:
        """mode 1: signal, 2: wait (for the epoch of the last signal), 3: barrier."""
        if self.device_epochs:
            self._check(self._lib.lsr_peer_sync_auto(self.rank, self.world, self._headers, mode, stream))
            return
        if mode & 1:
            self._barriers += 1
        fn = {1: self._lib.lsr_peer_signal, 2: self._lib.lsr_peer_wait, 3: self._lib.lsr_peer_barrier}[mode]
        self._check(fn(self.rank, self.world, self._headers, self._barriers & 0xFFFFFFFF, stream))

    def begin_step(self, slot: Optional[int] = None) -> "PeerFactorExchange":
        """Starts a step: the next `views_per_rank` backward passes write into this step's slot.  Returns the sink.
        Raises if a barrier of an earlier step timed out (a peer did not arrive within the time-out, LSR_PEER_TIMEOUT_MS
        / lsr_peer_set_timeout_ms, default 120 s): that step's sums were overwritten with NaN, never returned as
        gradients."""
        if self._flag_host.view(torch.int32)[0].item() != 0:
            raise self._err("PeerFactorExchange: a peer barrier timed out in an earlier step (a rank did not arrive); "
                            "that step's gradients were poisoned with NaN")
        self.epoch += 1
        self._count = 0
        self._slot = (self.epoch & 1) if slot is None else (int(slot) & 1)  # consecutive steps must alternate
        return self

    def new_factor(self, P: int, device, campos: torch.Tensor, degree: int) -> torch.Tensor:
        if P != self.P or self._count >= self.V:
            raise self._err(f"PeerFactorExchange: built for {self.V} view(s) of {self.P} Gaussians per step")
        i, slot = self._count, self._payload[self._slot]
        self._count += 1
        self._degree = int(degree)
        slot[self.cam_off + 4 * i: self.cam_off + 4 * i + 3].copy_(campos.reshape(3), non_blocking=True)
        return slot[i * P * 3:(i + 1) * P * 3].view(P, 3)

    def new_flat(self, n_floats: int, device) -> Optional[torch.Tensor]:
        """Memory for the parameter gradients of the step's FIRST backward: the head of this step's staging area, so
        that `exchange` finds them in place (autograd hands the slices to .grad and later views accumulate into them).
        None (caller allocates) for further views or when it does not fit."""
        if self.world == 1 or self._count != 0 or n_floats > self.grad_cap:  # asked before the view's new_factor;
            # grad_cap is 0 unless the exchange was built with grad_floats_per_gaussian (the all-reduce option)
            return None
        return self._payload[self._slot][self.in_off:self.in_off + n_floats]

    def finish(self, means3D: torch.Tensor, n_coeffs: int) -> torch.Tensor:
        """dL_dsh [P, n_coeffs, 3] summed over the views of all ranks; call once per step, after the backward passes.
        Every rank must have run the same number (1..views_per_rank) of backward passes since begin_step()."""
        return self.exchange(means3D, n_coeffs, [])[0]

    def exchange(self, means3D: torch.Tensor, n_coeffs: int, grads: Sequence[Optional[torch.Tensor]]):
        """The whole exchange of a step, all of it over peer memory: returns (dL_dsh summed over the views of all
        ranks, the other gradients summed over the ranks, in the order given; None stays None).
        The other gradients take a two-shot all-reduce (lsr_peer_allreduce): every rank reduces its 1/world chunk out
        of all ranks' staging areas and stores the sums into every rank's result area; the tensors returned are views
        of that area — valid until the step after next, i.e. consume them (optimizer step) within the step, as a
        training loop does.  Per rank: (gradients already sit in the staging area when the backward got its memory from
        new_flat, else one copy), barrier, then the all-reduce kernel + signal on the caller's stream while the
        SH-gradient kernel pulls the factors on a second stream, join, wait."""
        if self._count == 0:
            raise self._err("PeerFactorExchange: no view ran a backward this step")
        L, P, V = self._L, self.P, self._count
        stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        parity = self._slot
        slot = self._payload[parity]
        off = L.PEER_HEADER_BYTES + parity * self.slot_bytes
        # ---- stage the other gradients (adjacent slices travel as one copy) ----
        live = [g for g in grads if g is not None]
        places, total = {}, 0
        if live and self.world > 1:
            for g in live:
                if g.dtype != torch.float32 or not g.is_cuda:
                    raise self._err("PeerFactorExchange: gradients must be CUDA float32 tensors")
            fixed = [g if g.is_contiguous() else g.contiguous() for g in live]
            pieces = coalesce(fixed)
            stage0 = slot[self.in_off:self.in_off + 1].data_ptr()
            in_place = len(pieces) == 1 and pieces[0].data_ptr() == stage0  # new_flat's memory, handed back by autograd
            if not in_place:  # anything that happens to live in the staging area must move out of the way first
                lo, hi = stage0, stage0 + self.grad_cap * 4
                pieces = [p_.clone() if lo <= p_.data_ptr() < hi else p_ for p_ in pieces]
            for piece, orig in zip(pieces, coalesce(fixed)):
                n = piece.numel()
                if total + n > self.grad_cap:
                    raise self._err(f"PeerFactorExchange: more than {self.grad_cap} gradient floats per rank; raise "
                                    "grad_floats_per_gaussian")
                if not in_place:
                    slot[self.in_off + total: self.in_off + total + n].copy_(piece.reshape(-1), non_blocking=True)
                for g, f in zip(live, fixed):
                    d = f.data_ptr() - orig.data_ptr()
                    if 0 <= d < n * 4 and id(g) not in places:
                        places[id(g)] = total + d // 4
                total += (n + 3) // 4 * 4
        n = self.world * V
        g_ptrs, c_ptrs = (ctypes.c_void_p * n)(), (ctypes.c_void_p * n)()
        for r in range(self.world):  # rank-major, the order an all-gather would produce
            for i in range(V):
                g_ptrs[r * V + i] = self.bases[r] + off + i * P * 12
                c_ptrs[r * V + i] = self.bases[r] + off + (self.cam_off + 4 * i) * 4
        out = torch.empty((P, n_coeffs, 3), dtype=torch.float32, device=self.device)
        m3 = means3D.detach().contiguous()
        with torch.cuda.device(self.device):
            self._sync(3, stream)
            sh_stream = stream
            if total:  # the SH-gradient kernel runs beside the all-reduce kernel, on a second stream
                main = torch.cuda.current_stream(self.device)
                self._side.wait_stream(main)
                if not self.device_epochs:  # (under capture the join below already orders every later use)
                    out.record_stream(self._side)
                sh_stream = ctypes.c_void_p(self._side.cuda_stream)
            self._check(self._lib.lsr_sh_gradient_from_view_ptrs(P, n_coeffs, self._degree, n,
                                                                 ctypes.c_void_p(m3.data_ptr()), g_ptrs, c_ptrs,
                                                                 ctypes.c_void_p(out.data_ptr()), sh_stream))
            if total:
                ins = (ctypes.c_void_p * self.world)(*[b + off + self.in_off * 4 for b in self.bases])
                outs = (ctypes.c_void_p * self.world)(*[b + off + self.out_off * 4 for b in self.bases])
                self._check(self._lib.lsr_peer_allreduce(self.rank, self.world, ins, outs, total, stream))
                self._sync(1, stream)
                main.wait_stream(self._side)
                self._sync(2, stream)
            # a barrier that gave up leaves incomplete sums behind: never let them pass for gradients
            own = ctypes.c_void_p(self.own)
            self._check(self._lib.lsr_peer_poison_on_timeout(own, ctypes.c_void_p(out.data_ptr()), out.numel(),
                                                             stream))
            if total:
                self._check(self._lib.lsr_peer_poison_on_timeout(
                    own, ctypes.c_void_p(self.bases[self.rank] + off + self.out_off * 4), total, stream))
            self._flag_host.copy_(self._flag_dev, non_blocking=True)
        if not total:
            return out, list(grads)
        reduced = []
        for g in grads:
            if g is None:
                reduced.append(None)
            else:
                o = self.out_off + places[id(g)]
                reduced.append(slot[o:o + g.numel()].view(g.shape))
        return out, reduced

    def check(self) -> None:
        """Synchronises and raises if a barrier of this rank ever timed out (a peer was missing)."""
        with torch.cuda.device(self.device):
            self._check(self._lib.lsr_peer_status(ctypes.c_void_p(self.own), ctypes.c_void_p(
                torch.cuda.current_stream(self.device).cuda_stream)))

    def close(self) -> None:
        """Collective: every rank unmaps its peers, then (after a host barrier) frees its own buffer."""
        if self.own is None:
            return
        with torch.cuda.device(self.device):
            torch.cuda.synchronize(self.device)
            for p in self._opened:
                self._lib.lsr_peer_close(ctypes.c_void_p(p))
            self._opened = []
            if self.world > 1:
                dist.barrier(group=self.group)
            self._payload = None
            self._lib.lsr_peer_free(ctypes.c_void_p(self.own))
        self.own = None


def make_peer_exchange(P: int, views_per_rank: int, device, group=None, grad_floats_per_gaussian: int = 0,
                       device_epochs: bool = False) -> Optional[PeerFactorExchange]:
    """PeerFactorExchange if every rank could map every other rank's buffer, else None on all ranks (no NVLink /
    IPC between the processes): the caller then uses the all-gather path (`finish_sh_gradients`)."""
    ex, ok = None, 1
    try:
        ex = PeerFactorExchange(P, views_per_rank, device, group, grad_floats_per_gaussian, device_epochs)
    except Exception as e:  # noqa: BLE001 — whatever failed, the decision must be collective
        ok, ex = 0, None
        print(f"[langsplat_b200.dist] peer exchange unavailable on this rank: {e}", file=sys.stderr)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        flag = torch.tensor([ok], dtype=torch.int32, device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if int(flag.item()) == 0 and ex is not None:
            ex.own and ex._lib.lsr_peer_free(ctypes.c_void_p(ex.own))
            ex = None
    return ex
