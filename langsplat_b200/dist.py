"""View-sharded data parallelism (BASELINE.json config 4).

The reference is single-GPU (one camera per step, train.py:84-92).  The natural shard of the rasterizer is the camera
view: Gaussians are replicated, rank r of n renders views {r, r+n, ...} of the step and runs its own backward; the one
exchange step is a sum-allreduce of the Gaussian-parameter gradients (NCCL over NVLink/NVSwitch on the GPU box, gloo
in the CPU tests).  No collective touches the rasterizer's data path itself.
"""
from __future__ import annotations

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
