"""Fused Adam for the Gaussian parameters.

`FusedAdam` takes the place of `torch.optim.Adam(l, lr=0.0, eps=1e-15)` in GaussianModel.training_setup
(scene/gaussian_model.py:198-227): same constructor, same `param_groups` (per-group `lr`, `name`, …), same per-parameter
state keys (`step`, `exp_avg`, `exp_avg_sq`), so the reference's densification code that edits optimizer state
(cat_tensors_to_optimizer / _prune_optimizer / replace_tensor_to_optimizer, scene/gaussian_model.py:300-372) and
state_dict round trips with torch.optim.Adam work unchanged.  `step()` updates every group in ONE kernel launch
(csrc/adam.cu) with PyTorch's single-tensor Adam arithmetic.  `step(visible=mask)` restricts the update — moments
included — to the Gaussians with mask set (the later 3DGS "sparse Adam"); the default reproduces the reference.
CUDA float32 parameters only; no CPU path.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from . import _lib as L
from .rasterizer import LsrError


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8):
        if lr < 0.0 or eps < 0.0 or not (0.0 <= betas[0] < 1.0) or not (0.0 <= betas[1] < 1.0):
            raise ValueError(f"invalid Adam hyper-parameters: lr={lr} betas={betas} eps={eps}")
        # the keys torch.optim.Adam puts in a group, so that state_dicts are interchangeable
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False, maximize=False, foreach=None,
                        capturable=False, differentiable=False, fused=None)
        super().__init__(params, defaults)

    @torch.no_grad()
    def step(self, closure=None, visible: Optional[torch.Tensor] = None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        batches = {}  # (device, step, betas, eps) -> list of lsr_adam_tensor fields
        keep = []
        for group in self.param_groups:
            if group.get("weight_decay", 0) or group.get("amsgrad", False) or group.get("maximize", False):
                raise LsrError("FusedAdam: weight_decay / amsgrad / maximize are not supported (the reference uses none)")
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise LsrError("FusedAdam: parameters must be contiguous CUDA float32 tensors (no CPU path)")
                if p.grad.is_sparse:
                    raise LsrError("FusedAdam: sparse gradients are not supported")
                state = self.state[p]
                if len(state) == 0:  # torch/optim/adam.py _init_group
                    state["step"] = torch.tensor(0.0, dtype=torch.float32)
                    state["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    state["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                state["step"] += 1
                grad = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                m, v = state["exp_avg"], state["exp_avg_sq"]
                if not (m.is_contiguous() and v.is_contiguous()):
                    raise LsrError("FusedAdam: optimizer state must be contiguous")
                keep.append(grad)
                row = p.numel() // p.shape[0] if p.dim() > 0 and p.shape[0] > 0 else 1
                if visible is not None and p.dim() > 0 and p.shape[0] != visible.shape[0]:
                    raise LsrError("FusedAdam: `visible` must have one entry per Gaussian (first dimension)")
                key = (p.device, int(state["step"].item()), float(beta1), float(beta2), float(group["eps"]))
                batches.setdefault(key, []).append((p.data_ptr(), grad.data_ptr(), m.data_ptr(), v.data_ptr(),
                                                    p.numel(), row, float(group["lr"])))
        vis = None
        if visible is not None:
            if not visible.is_cuda:
                raise LsrError("FusedAdam: `visible` must be a CUDA tensor")
            vis = (visible if visible.dtype in (torch.bool, torch.uint8) else visible != 0).contiguous()
        lib = L.load()
        for (dev, step, beta1, beta2, eps), items in batches.items():
            with torch.cuda.device(dev):
                stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
                for i in range(0, len(items), L.ADAM_MAX_TENSORS):
                    chunk = items[i:i + L.ADAM_MAX_TENSORS]
                    arr = (L.AdamTensor * len(chunk))(*[L.AdamTensor(*it) for it in chunk])
                    rc = lib.lsr_adam_step(len(chunk), arr, step, beta1, beta2, eps,
                                           None if vis is None else ctypes.c_void_p(vis.data_ptr()), stream)
                    if rc != L.LSR_OK:
                        raise LsrError(L.last_error())
        return loss
