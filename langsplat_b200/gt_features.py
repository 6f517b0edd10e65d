"""Ground-truth language-feature maps on the GPU, cached per camera.

`Camera.get_language_feature(language_feature_dir, feature_level)` (scene/cameras.py:58-92) reloads
`<image_name>_s.npy` / `<image_name>_f.npy`, builds a meshgrid and gathers `feature_map[seg]` on the CPU on every
iteration, then uploads the [C,H,W] result.  `get_language_feature(camera, language_feature_dir, feature_level)` here
returns the same `(point_feature [C,H,W], mask [1,H,W] bool)` pair, bit-identical, but loads the two files once per
camera, keeps the compact seg map and feature table on the GPU and gathers there (csrc/gt_features.cu); the assembled
maps are cached too unless `keep_maps=False`.  CUDA only; no CPU path.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from . import _lib as L
from .rasterizer import LsrError


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def language_feature_map(seg_map: torch.Tensor, feature_map: torch.Tensor, feature_level: int,
                         image_height: Optional[int] = None, image_width: Optional[int] = None,
                         validate: bool = True, float_mask: bool = False):
    """seg_map [L,Hs,Ws] (CUDA float32 or int32), feature_map [K,C] (CUDA float32) ->
    (point_feature [C,H,W] float32, mask [1,H,W] bool) of level `feature_level`, H x W = the top-left crop the
    reference's meshgrid takes (defaults: the whole map).  `float_mask=True` appends the mask as float32 (what
    langsplat_b200.l1_loss consumes).  `validate` raises IndexError for indices outside [-K, K) like PyTorch would."""
    if seg_map.dim() != 3 or feature_map.dim() != 2:
        raise LsrError("language_feature_map: seg_map must be [L,H,W] and feature_map [K,C]")
    if not (0 <= feature_level < seg_map.shape[0]) or feature_level > 3:
        raise ValueError("feature_level=", feature_level)  # scene/cameras.py:87-88
    if not seg_map.is_cuda or not feature_map.is_cuda:
        raise LsrError("language_feature_map: CUDA tensors required (langsplat_b200 has no CPU path)")
    if seg_map.dtype not in (torch.float32, torch.int32) or feature_map.dtype != torch.float32:
        raise LsrError("language_feature_map: seg_map must be float32 or int32, feature_map float32")
    H = seg_map.shape[1] if image_height is None else int(image_height)
    W = seg_map.shape[2] if image_width is None else int(image_width)
    if H > seg_map.shape[1] or W > seg_map.shape[2]:
        raise IndexError(f"seg map {tuple(seg_map.shape[1:])} is smaller than the image {H}x{W}")
    seg = seg_map[feature_level]
    if seg.stride(1) != 1:
        seg = seg.contiguous()
    table = feature_map.contiguous()
    K, C = table.shape
    dev = seg.device
    out = torch.empty((C, H, W), dtype=torch.float32, device=dev)
    mask = torch.empty((1, H, W), dtype=torch.bool, device=dev)
    mask_f = torch.empty((1, H, W), dtype=torch.float32, device=dev) if float_mask else None
    bad = torch.zeros((1,), dtype=torch.int32, device=dev) if validate else None
    lib = L.load()
    with torch.cuda.device(dev):
        rc = lib.lsr_language_feature_map(_ptr(seg), int(seg.dtype == torch.float32), seg.stride(0), W, H, _ptr(table),
                                          K, C, _ptr(out), _ptr(mask), _ptr(mask_f), _ptr(bad),
                                          ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    if rc != L.LSR_OK:
        raise LsrError(L.last_error())
    if validate and int(bad.item()) != 0:
        raise IndexError(f"{int(bad.item())} seg-map indices are out of range for a feature table of {K} rows")
    return (out, mask, mask_f) if float_mask else (out, mask)


class LanguageFeatureCache:
    """Per-camera cache: the seg map and feature table of each image live on the GPU after the first request, and
    (keep_maps) so do the assembled maps per level."""

    def __init__(self, device="cuda", keep_maps: bool = True):
        self.device, self.keep_maps = torch.device(device), keep_maps
        self._inputs: Dict[str, Tuple[torch.Tensor, torch.Tensor]] = {}
        self._maps: Dict[Tuple[str, int, int, int], Tuple[torch.Tensor, torch.Tensor]] = {}

    def inputs(self, language_feature_name: str):
        got = self._inputs.get(language_feature_name)
        if got is None:
            seg = np.load(language_feature_name + "_s.npy")  # scene/cameras.py:60-61
            feat = np.load(language_feature_name + "_f.npy")
            if seg.dtype not in (np.float32, np.int32):
                seg = seg.astype(np.int32) if np.issubdtype(seg.dtype, np.integer) else seg.astype(np.float32)
            got = (torch.from_numpy(seg).to(self.device), torch.from_numpy(feat.astype(np.float32)).to(self.device))
            self._inputs[language_feature_name] = got
        return got

    def get(self, language_feature_name: str, feature_level: int, image_height: int, image_width: int):
        key = (language_feature_name, int(feature_level), int(image_height), int(image_width))
        got = self._maps.get(key)
        if got is None:
            seg, table = self.inputs(language_feature_name)
            got = language_feature_map(seg, table, feature_level, image_height, image_width)
            if self.keep_maps:
                self._maps[key] = got
        return got

    def clear(self) -> None:
        self._inputs.clear()
        self._maps.clear()


_default_cache: Optional[LanguageFeatureCache] = None


def get_language_feature(camera, language_feature_dir: str, feature_level: int,
                         cache: Optional[LanguageFeatureCache] = None):
    """Drop-in for `camera.get_language_feature(language_feature_dir, feature_level)` (scene/cameras.py:58-92); `camera`
    needs image_name, image_height and image_width.  Returns (point_feature [C,H,W], mask [1,H,W] bool) on the GPU."""
    global _default_cache
    if cache is None:
        if _default_cache is None:
            _default_cache = LanguageFeatureCache()
        cache = _default_cache
    name = os.path.join(language_feature_dir, camera.image_name)
    return cache.get(name, feature_level, camera.image_height, camera.image_width)
