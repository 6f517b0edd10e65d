"""langsplat_b200 — B200-native (sm_100a) differentiable Gaussian-splatting rasterizer.

Drop-in for LangSplat's `diff_gaussian_rasterization` extension: same `GaussianRasterizationSettings` /
`GaussianRasterizer` surface, hand-written CUDA kernels behind a C ABI (include/lsr_api.h -> liblsr.so).
Importing this package does not load the CUDA library; the first rasterizer call does, and fails loudly if
liblsr.so is missing (there is no CPU fallback).
"""
from .rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, LsrError, rasterize_gaussians,
                         cpu_deep_copy_tuple, set_async_forward, async_forward, check_async_status,
                         reserve_instances)
from .loss import l1_loss
from .activations import activate

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians", "LsrError",
           "cpu_deep_copy_tuple", "l1_loss", "activate", "set_async_forward", "async_forward",
           "check_async_status", "reserve_instances"]
