"""Eval-side decode of rendered language features.

The reference restores 512-d CLIP features from the 3-d rendered ones with the autoencoder's decoder
(`Autoencoder.decode`, autoencoder/model.py:42-46; call site eval/evaluate_iou_loc.py:258-261) and then scores every
pixel against the phrase embeddings (`OpenCLIPNetwork.get_relevancy` / `get_max_across`,
eval/openclip_encoder.py:42-56, 96-111).  `FeatureDecoder` mirrors those entry points over ONE fused CUDA kernel
(csrc/decoder.cu, `lsr_decode`): `decode(x)` returns the unit 512-d features, `get_max_across(sem_map3)` goes from the
3-d maps straight to the relevancy maps without ever materialising the 512-d features in HBM.

Precisions, all with fp32 accumulation (in TMEM for the tensor-core variants):
  "bf16x2" (default) — tcgen05 tensor cores at the reference's arithmetic: the reference runs fp32 SGEMMs, so every fp32
      operand is carried as two bf16 numbers (hi + lo, 16 mantissa bits) and every product as three bf16 MMAs; results
      agree with the fp32 chain to ~1e-6 (the bars of the fp32 variant);
  "tf32" / "f16" — one MMA per product, operands rounded to 10 mantissa bits (TF32, or fp16 clamped to +-65504: half the
      bytes, twice the MMA rate, two CTAs per SM): faster, errors ~1e-4;
  "fp32" — IEEE fp32 on the FP32 pipe.
CUDA only; no CPU path.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence

import torch

from . import _lib as L
from .rasterizer import LsrError

_PRECISIONS = {"fp32": L.DECODE_FP32, "tf32": L.DECODE_TF32, "f16": L.DECODE_F16, "bf16x2": L.DECODE_BF16X2}


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


class FeatureDecoder:
    """Decoder MLP of the reference's Autoencoder plus the phrase relevancy, fused.

    weights[l]: nn.Linear weight [out, in]; biases[l]: [out] (CUDA float32).  `from_state_dict` takes the checkpoint the
    reference saves (autoencoder/train.py: keys `decoder.<i>.weight/bias`, ReLU modules interleaved)."""

    def __init__(self, weights: Sequence[torch.Tensor], biases: Sequence[torch.Tensor], precision: str = "bf16x2"):
        if len(weights) != len(biases) or not 1 <= len(weights) <= L.DECODER_MAX_LAYERS:
            raise LsrError(f"FeatureDecoder: 1..{L.DECODER_MAX_LAYERS} (weight, bias) pairs expected")
        if precision not in _PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(_PRECISIONS)}")
        self.weights = [w.detach().to(torch.float32).contiguous() for w in weights]
        self.biases = [b.detach().to(torch.float32).contiguous() for b in biases]
        for w, b in zip(self.weights, self.biases):
            if not (w.is_cuda and b.is_cuda):
                raise LsrError("FeatureDecoder: CUDA tensors required (langsplat_b200 has no CPU path)")
            if w.dim() != 2 or b.shape != (w.shape[0],):
                raise LsrError("FeatureDecoder: weight [out, in] and bias [out] expected")
        for a, b in zip(self.weights[:-1], self.weights[1:]):
            if b.shape[1] != a.shape[0]:
                raise LsrError("FeatureDecoder: layer widths do not chain")
        self.device = self.weights[0].device
        self.precision = precision
        self.in_dim, self.out_dim = self.weights[0].shape[1], self.weights[-1].shape[0]
        d = L.Decoder()
        d.n_layers = len(self.weights)
        d.dims[0] = self.in_dim
        for l, (w, b) in enumerate(zip(self.weights, self.biases)):
            d.dims[l + 1] = w.shape[0]
            d.weight[l] = w.data_ptr()
            d.bias[l] = b.data_ptr()
        d.packed_tf32 = d.packed_f16 = d.packed_bf16x2 = None
        self._desc = d
        self._packed = {}
        self.pos_embeds = self.neg_embeds = self._phrases = None

    @classmethod
    def from_state_dict(cls, state_dict, device="cuda", precision: str = "bf16x2"):
        idx = sorted({int(k.split(".")[1]) for k in state_dict if k.startswith("decoder.") and k.endswith(".weight")})
        if not idx:
            raise LsrError("no decoder.<i>.weight entries in the state dict")
        ws = [state_dict[f"decoder.{i}.weight"].to(device) for i in idx]
        bs = [state_dict[f"decoder.{i}.bias"].to(device) for i in idx]
        return cls(ws, bs, precision)

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _ensure_packed(self, lib, prec):
        if prec not in self._packed:
            nbytes = lib.lsr_decoder_packed_bytes(ctypes.byref(self._desc), prec)
            if nbytes == 0:
                raise LsrError(L.last_error())
            packed = torch.empty(nbytes // 4, dtype=torch.float32, device=self.device)
            rc = lib.lsr_decoder_pack(ctypes.byref(self._desc), prec, _ptr(packed), self._stream())
            if rc != L.LSR_OK:
                raise LsrError(L.last_error())
            self._packed[prec] = packed
            if prec == L.DECODE_F16:
                self._desc.packed_f16 = packed.data_ptr()
            elif prec == L.DECODE_BF16X2:
                self._desc.packed_bf16x2 = packed.data_ptr()
            else:
                self._desc.packed_tf32 = packed.data_ptr()

    def _run(self, x: torch.Tensor, want_out: bool, want_relevancy: bool, precision: Optional[str]):
        prec = _PRECISIONS[precision or self.precision]
        if not x.is_cuda or x.dtype != torch.float32:
            raise LsrError("FeatureDecoder: CUDA float32 input required")
        if x.shape[-1] != self.in_dim:
            raise LsrError(f"FeatureDecoder: last dimension must be {self.in_dim}")
        x2 = x.reshape(-1, self.in_dim).contiguous()
        n = x2.shape[0]
        out = torch.empty((n, self.out_dim), dtype=torch.float32, device=self.device) if want_out else None
        rel = None
        n_pos = n_neg = 0
        if want_relevancy:
            if self._phrases is None:
                raise LsrError("FeatureDecoder: set_phrases(pos_embeds, neg_embeds) first")
            n_pos, n_neg = self.pos_embeds.shape[0], self.neg_embeds.shape[0]
            rel = torch.empty((n_pos, n), dtype=torch.float32, device=self.device)
        lib = L.load()
        if n == 0:
            return out, rel
        with torch.cuda.device(self.device):
            if prec != L.DECODE_FP32:
                self._ensure_packed(lib, prec)
            work = None
            if want_relevancy and prec != L.DECODE_FP32:  # per call: safe with several streams in flight
                work = torch.empty(lib.lsr_decode_work_bytes() // 4, dtype=torch.float32, device=self.device)
            rc = lib.lsr_decode(ctypes.byref(self._desc), _ptr(x2), n, prec, _ptr(out), _ptr(self._phrases), n_pos,
                                n_neg, _ptr(rel), _ptr(work), 0 if work is None else work.numel() * 4, self._stream())
        if rc != L.LSR_OK:
            raise LsrError(L.last_error())
        return out, rel

    # ---- Autoencoder.decode ----
    def decode(self, x: torch.Tensor, precision: Optional[str] = None) -> torch.Tensor:
        """x [..., 3] -> [..., 512] unit vectors (autoencoder/model.py:42-46)."""
        out, _ = self._run(x, True, False, precision)
        return out.view(*x.shape[:-1], self.out_dim)

    # ---- OpenCLIPNetwork ----
    def set_phrases(self, pos_embeds: torch.Tensor, neg_embeds: torch.Tensor):
        """Unit phrase embeddings as OpenCLIPNetwork keeps them (eval/openclip_encoder.py:33-40, 64-72)."""
        pos = pos_embeds.detach().to(self.device, torch.float32).reshape(-1, self.out_dim)
        neg = neg_embeds.detach().to(self.device, torch.float32).reshape(-1, self.out_dim)
        if pos.shape[0] < 1 or neg.shape[0] < 1 or pos.shape[0] + neg.shape[0] > L.DECODER_MAX_PHRASES:
            raise LsrError(f"set_phrases: at least one positive and one negative, at most {L.DECODER_MAX_PHRASES}")
        self.pos_embeds, self.neg_embeds = pos, neg
        self._phrases = torch.cat([pos, neg], 0).contiguous()

    def relevancy_maps(self, x: torch.Tensor, precision: Optional[str] = None) -> torch.Tensor:
        """x [..., 3] -> [n_pos, ...]: the positive probability get_relevancy(decode(x), j)[..., 0] for every j."""
        _, rel = self._run(x, False, True, precision)
        return rel.view(rel.shape[0], *x.shape[:-1])

    def get_relevancy(self, x: torch.Tensor, positive_id: int, precision: Optional[str] = None) -> torch.Tensor:
        """get_relevancy(embed, positive_id) of the reference with embed = decode(x): [N, 2] (positive, negative)
        probabilities of the pair with the lowest positive probability (eval/openclip_encoder.py:42-56)."""
        p = self.relevancy_maps(x.reshape(-1, self.in_dim), precision)[positive_id]
        return torch.stack((p, 1.0 - p), -1)

    def get_max_across(self, sem_map: torch.Tensor, precision: Optional[str] = None) -> torch.Tensor:
        """sem_map [levels, H, W, 3] (the 3-d maps, not the restored ones) -> relev_map [levels, n_phrases, H, W]
        (eval/openclip_encoder.py:96-111)."""
        return self.relevancy_maps(sem_map, precision).permute(1, 0, 2, 3).contiguous()
