"""Fused (masked) L1 loss on the rendered images — the step that follows the rasterizer in LangSplat's training loop.

Mirrors `l1_loss(network_output, gt)` of utils/loss_utils.py:17-18; with `mask` it computes what train.py:96-99 writes
as `l1_loss(language_feature * language_feature_mask, gt_language_feature * language_feature_mask)` without the
two masked temporaries.  One CUDA kernel forward and one backward (csrc/loss.cu) instead of PyTorch's nine elementwise /
reduction launches.  CUDA float32 only; no CPU path.

`gt` may also be the uint8 image as it is stored on disk ([C,H,W], values 0..255): the kernels then form
`gt = u / 255` themselves — the value utils/general_utils.py:23 (PILtoTorch) computes on the host — so a training loop
uploads 1 byte per element per view instead of 4 and no float copy of the image is ever written to device memory.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from . import _lib as L
from .rasterizer import LsrError


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _check(t: torch.Tensor, name: str) -> torch.Tensor:
    if not t.is_cuda or t.dtype != torch.float32:
        raise LsrError(f"l1_loss: {name} must be a CUDA float32 tensor (langsplat_b200 has no CPU path)")
    return t.contiguous()


def _mask_rows(mask: Optional[torch.Tensor], pred: torch.Tensor) -> Optional[torch.Tensor]:
    """The mask is either broadcast over the leading (channel) dimension ([1,H,W] / [H,W], as
    scene/cameras.py:58-92 returns it) or has pred's full shape (then it is applied as one [1, C*N] row)."""
    if mask is None:
        return None
    if mask.is_cuda and mask.dtype in (torch.bool, torch.uint8):
        # Camera.get_language_feature returns `seg != -1`, a bool mask (scene/cameras.py:70), and train.py:96-99
        # multiplies it in as it is; the kernels read a float mask
        mask = mask.to(torch.float32)
    mask = _check(mask, "mask")
    if mask.numel() == pred.numel():
        return mask
    if pred.dim() >= 1 and mask.numel() * pred.shape[0] == pred.numel():
        return mask
    raise LsrError(f"l1_loss: mask of shape {tuple(mask.shape)} does not broadcast over {tuple(pred.shape)}")


class _L1Loss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, gt, mask):
        pred_c = _check(pred, "network_output")
        u8 = gt.dtype == torch.uint8
        if u8:
            if not gt.is_cuda:
                raise LsrError("l1_loss: gt must be a CUDA tensor (langsplat_b200 has no CPU path)")
            gt_c = gt.contiguous()
        else:
            gt_c = _check(gt, "gt")
        if pred_c.shape != gt_c.shape:
            raise LsrError(f"l1_loss: shapes differ: {tuple(pred_c.shape)} vs {tuple(gt_c.shape)}")
        m = _mask_rows(mask, pred_c)
        total = pred_c.numel()
        if m is None or m.numel() == total:
            C, N = 1, total
        else:
            C, N = pred_c.shape[0], total // pred_c.shape[0]
        lib = L.load()
        loss = torch.empty((), dtype=torch.float32, device=pred_c.device)
        stream = ctypes.c_void_p(torch.cuda.current_stream(pred_c.device).cuda_stream)
        with torch.cuda.device(pred_c.device):
            fwd = lib.lsr_l1_loss_forward_u8 if u8 else lib.lsr_l1_loss_forward
            rc = fwd(_ptr(pred_c), _ptr(gt_c), _ptr(m), C, N, _ptr(loss), stream)
        if rc != L.LSR_OK:
            raise LsrError(L.last_error())
        if total == 0:
            loss.fill_(float("nan"))  # torch: mean of an empty tensor
        ctx.save_for_backward(pred_c, gt_c, m if m is not None else torch.empty(0, device=pred_c.device))
        ctx.dims = (C, N, m is not None, u8)
        return loss

    @staticmethod
    def backward(ctx, grad_out):
        pred_c, gt_c, m = ctx.saved_tensors
        C, N, has_mask, u8 = ctx.dims
        if not ctx.needs_input_grad[0]:
            return None, None, None
        lib = L.load()
        g = grad_out.to(dtype=torch.float32, device=pred_c.device).contiguous()
        out = torch.empty_like(pred_c)
        stream = ctypes.c_void_p(torch.cuda.current_stream(pred_c.device).cuda_stream)
        with torch.cuda.device(pred_c.device):
            bwd = lib.lsr_l1_loss_backward_u8 if u8 else lib.lsr_l1_loss_backward
            rc = bwd(_ptr(pred_c), _ptr(gt_c), _ptr(m) if has_mask else None, C, N, _ptr(g), _ptr(out), stream)
        if rc != L.LSR_OK:
            raise LsrError(L.last_error())
        return out, None, None


def l1_loss(network_output: torch.Tensor, gt: torch.Tensor, mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """`torch.abs(network_output * mask - gt * mask).mean()` (mask = ones when omitted), differentiable w.r.t.
    `network_output` only (the ground truth and the mask are data, as in train.py)."""
    return _L1Loss.apply(network_output, gt, mask)
