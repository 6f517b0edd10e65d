"""Fused per-Gaussian activations in front of the rasterizer.

`activate(scaling, rotation, opacity, language_feature)` returns what LangSplat's renderer feeds the rasterizer:
`pc.get_scaling` = exp (scene/gaussian_model.py:33,134-136), `pc.get_rotation` = F.normalize (:40,138-140),
`pc.get_opacity` = sigmoid (:37,153-155) and the normalised language feature `f / (f.norm(dim=-1, keepdim=True) + 1e-9)`
(gaussian_renderer/__init__.py:86-88) — one CUDA kernel forward and one backward (csrc/activations.cu) instead of
PyTorch's ten-odd launches and [P,·] temporaries.  Any argument may be None (it is skipped and None is returned in its
place).  CUDA float32 only; no CPU path.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch

from . import _lib as L
from .rasterizer import LsrError

_SHAPES = (("scaling", 3), ("rotation", 4), ("opacity", 1), ("language_feature", 3))


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _prepare(tensors) -> Tuple[list, int, torch.device]:
    P, dev, out = None, None, []
    for (name, width), t in zip(_SHAPES, tensors):
        if t is None:
            out.append(None)
            continue
        if not t.is_cuda or t.dtype != torch.float32:
            raise LsrError(f"activate: {name} must be a CUDA float32 tensor (langsplat_b200 has no CPU path)")
        if t.dim() != 2 or t.shape[1] != width:
            raise LsrError(f"activate: {name} must be [P,{width}], got {tuple(t.shape)}")
        if P is not None and t.shape[0] != P:
            raise LsrError(f"activate: {name} has {t.shape[0]} rows, expected {P}")
        P, dev = t.shape[0], t.device
        out.append(t.contiguous())
    if P is None:
        raise LsrError("activate: nothing to do (every argument is None)")
    return out, P, dev


class _Activate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, scaling, rotation, opacity, language_feature):
        raw, P, dev = _prepare((scaling, rotation, opacity, language_feature))
        outs = [None if t is None else torch.empty_like(t) for t in raw]
        lib = L.load()
        with torch.cuda.device(dev):
            rc = lib.lsr_activate_forward(P, *[_ptr(t) for t in raw], *[_ptr(t) for t in outs],
                                          ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
        if rc != L.LSR_OK:
            raise LsrError(L.last_error())
        ctx.save_for_backward(*[t for t in raw if t is not None])
        ctx.present = [t is not None for t in raw]
        ctx.P, ctx.dev = P, dev
        ctx.set_materialize_grads(False)
        return tuple(outs)

    @staticmethod
    def backward(ctx, *grads):
        saved = list(ctx.saved_tensors)
        raw = [saved.pop(0) if p else None for p in ctx.present]
        ups, outs = [], []
        for i, (t, g) in enumerate(zip(raw, grads)):
            if t is None or g is None or not ctx.needs_input_grad[i]:
                ups.append(None)
                outs.append(None)
                continue
            ups.append(g.contiguous())
            outs.append(torch.empty_like(t))
        if any(o is not None for o in outs):
            lib = L.load()
            with torch.cuda.device(ctx.dev):
                rc = lib.lsr_activate_backward(ctx.P, *[_ptr(t) for t in raw], *[_ptr(t) for t in ups],
                                               *[_ptr(t) for t in outs],
                                               ctypes.c_void_p(torch.cuda.current_stream(ctx.dev).cuda_stream))
            if rc != L.LSR_OK:
                raise LsrError(L.last_error())
        return tuple(outs)


def activate(scaling: Optional[torch.Tensor] = None, rotation: Optional[torch.Tensor] = None,
             opacity: Optional[torch.Tensor] = None, language_feature: Optional[torch.Tensor] = None):
    """(scales, rotations, opacities, language_feature_precomp) for the rasterizer from the raw parameters."""
    return _Activate.apply(scaling, rotation, opacity, language_feature)
