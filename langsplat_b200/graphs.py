"""CUDA-graph capture of a fixed-shape rasterizer step.

The reference cannot be captured: its forward reads the instance count back to the host in the middle of every frame
(RAST/cuda_rasterizer/rasterizer_impl.cu:284), and a synchronising call is illegal inside a stream capture.  This
package's asynchronous forward (lsr_forward_nosync, `set_async_forward`) makes no such call, so a whole training
iteration — forward, loss, backward — can be recorded once and replayed with one launch: the per-iteration host work
(Python, ctypes marshalling, ~20 kernel launches, allocator calls) shrinks to a `replay()`.

What changes from step to step must live in tensors that keep their address: the camera matrices (`viewmatrix`,
`projmatrix`, `campos` of the GaussianRasterizationSettings), the ground truth, the parameters.  Image size, field of
view, SH degree and the Gaussian count are baked into the capture.  The instance list is sized while capturing (from the
frames the device has seen, times `capacity_margin`, see `reserve_instances`); a replayed frame that outgrows it has its
images NaN-filled on the device — call `GraphedStep.check()` (or look at the loss) to find out.
"""
from __future__ import annotations

from typing import Any, Callable, Optional

import torch

from . import _lib as L
from . import rasterizer as R


class GraphedStep:
    """`fn()` runs one iteration on static tensors and returns tensors (or a tuple / dict of them) that the caller
    reads after each `replay()`.  Gradients that `fn` leaves in `.grad` attributes keep their addresses too.

        step = GraphedStep(fn)          # warms up, then captures
        view.copy_(new_view); gt.copy_(new_gt)
        out = step.replay()             # the tensors `fn` returned, refreshed
    """

    def __init__(self, fn: Callable[[], Any], warmup: int = 3, capacity_margin: Optional[float] = None,
                 pool=None, capture_error_mode: str = "global"):
        """capture_error_mode: torch.cuda.graph's; "thread_local" when other threads of the process may call CUDA
        while this one captures (torch.distributed's NCCL watchdog polls its events)."""
        if not torch.cuda.is_available():
            raise R.LsrError("GraphedStep needs a CUDA device (langsplat_b200 has no CPU path)")
        with R.async_forward(True, capacity_margin):
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(max(1, warmup)):  # allocations, lazy initialisation, and the device learns the count
                    fn()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            R.check_async_status()
            before = {k: len(v) for k, v in R._pending.items()}
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph, pool=pool, capture_error_mode=capture_error_mode):
                self.outputs = fn()
        # the forwards recorded by the capture: their status words are refreshed by every replay (same buffers), but
        # the stamp protocol cannot tell replays apart, so they leave the pending list and are read by check()
        self._frames = []
        for dev_idx, pend in R._pending.items():
            self._frames.extend(pend[before.get(dev_idx, 0):])
            del pend[before.get(dev_idx, 0):]

    def replay(self):
        self.graph.replay()
        return self.outputs

    def check(self) -> None:
        """Waits for the device and raises if a replayed frame outgrew its instance list (its images were NaN-filled)
        or tripped the prefiltered trap."""
        torch.cuda.synchronize()
        for st in self._frames:  # rasterizer._AsyncFrame records; their pinned status words are rewritten by every replay
            words = st.status[1] if st.status is not None else None
            if words is None:
                continue
            flags, count = int(words[L.STATUS_FLAGS]), int(words[L.STATUS_NUM_RENDERED])
            if flags & L.STATUS_FLAG_OVERFLOW:
                raise R.LsrError(f"graphed step: a frame has {count} Gaussian/tile instances but the captured instance "
                                 f"list holds {st.capacity}; its images were filled with NaN.  Capture again after "
                                 "reserve_instances(device, count) or with a larger capacity_margin.")
            if flags & L.STATUS_FLAG_PREFILTERED:
                raise R.LsrError("Point is filtered although prefiltered is set. This shouldn't happen!")
