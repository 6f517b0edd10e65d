"""View-sharded data parallelism (BASELINE.json config 4).

The reference is single-GPU (one camera per step, train.py:84-92).  The natural shard of the rasterizer is the camera
view: Gaussians are replicated, rank r of n renders views {r, r+n, ...} of the step and runs its own backward; the one
exchange step is a sum-allreduce of the Gaussian-parameter gradients (NCCL over NVLink/NVSwitch on the GPU box, gloo
in the CPU tests).  No collective touches the rasterizer's data path itself.
"""
from __future__ import annotations

import contextlib
import ctypes
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
def factored_sh_gradients():
    """Context in which GaussianRasterizer backward passes collect per-view SH-gradient factors (yields the list)."""
    from . import rasterizer as R
    sink: List = []
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
