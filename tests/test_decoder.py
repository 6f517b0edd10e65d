"""Eval-side decode: decoder MLP + phrase relevancy.  The oracle's float64 restatement is pinned
on vectors produced by the reference's own Autoencoder.decode and OpenCLIPNetwork.get_relevancy
(tests/golden/make_golden_decoder.py); on the GPU langsplat_b200.decode.FeatureDecoder (lsr_decode through the C ABI)
is held to the same vectors and to the oracle at ragged sizes.
Tolerances (outputs are unit 512-vectors, components ~0.04, and probabilities):
  fp32 variant: 1e-5 on features, 2e-5 on relevancy (summation order only);
  tf32 variant (tcgen05, TF32 operands, fp32 accumulate): 5e-4 on features, 1e-3 on relevancy — TF32 operand truncation
  emulated on the CPU for the golden case gives 1.1e-4 / 2.4e-4, and the B200 reproduces exactly that;
  f16 variant (operands rounded to nearest fp16, same 10 mantissa bits): same bars (emulation: 3.5e-5 / 6e-5)."""
import os

import numpy as np
import pytest
import torch

from helpers import orc

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "decoder.npz"))
NL = int(GOLD["n_layers"])
WS = [GOLD[f"w{i}"] for i in range(NL)]
BS = [GOLD[f"b{i}"] for i in range(NL)]
N_POS = int(GOLD["n_pos"])
# (features, relevancy) max-abs bars.  "bf16x2" — the tensor-core variant at the reference's arithmetic (split-bf16
# operands, three MMAs per product) — is held to the bars of the IEEE-fp32 variant; the single-MMA variants carry the
# error of 10-bit operands.
TOL = {"fp32": (1e-5, 2e-5), "bf16x2": (1e-5, 2e-5), "tf32": (5e-4, 1e-3), "f16": (5e-4, 1e-3)}


def test_oracle_matches_reference_vectors():
    y = orc.decode_features(GOLD["x"], WS, BS)
    assert np.abs(y - GOLD["y"]).max() < 5e-7
    r = orc.relevancy(GOLD["y"], GOLD["phrases"], N_POS)
    assert np.abs(r - GOLD["relevancy"]).max() < 1e-6
    assert np.abs(r.sum(-1) - 1).max() < 1e-12


def test_oracle_decode_is_scale_free_after_first_relu_free_layer():
    # property used at full size on the GPU: outputs are unit vectors
    y = orc.decode_features(np.random.default_rng(0).normal(size=(64, 3)), WS, BS)
    assert np.abs(np.sqrt((y * y).sum(-1)) - 1).max() < 1e-12


def test_operand_rounding_explains_the_tensor_core_tolerance():
    """Where the 5e-4 / 1e-3 bars come from: rounding activations and weights to 10 mantissa bits before every
    product, emulated on the CPU, against the reference's vectors.  The B200 measured 1.09e-4 / 2.44e-4 with raw fp32
    operands (tensor-core truncation) and 3.46e-5 / 5.8e-5 with fp16 operands — the figures this emulation gives."""
    got = {}
    for operand in ("tf32_trunc", "tf32", "f16"):
        y = orc.decode_features(GOLD["x"], WS, BS, operand=operand)
        r = orc.relevancy(y, GOLD["phrases"], N_POS)[..., 0]
        got[operand] = (np.abs(y - GOLD["y"]).max(), np.abs(r - GOLD["relevancy"][..., 0]).max())
        assert got[operand][0] < TOL["tf32"][0] / 2 and got[operand][1] < TOL["tf32"][1] / 2
    assert 0.9e-4 < got["tf32_trunc"][0] < 1.3e-4 and 2.0e-4 < got["tf32_trunc"][1] < 2.9e-4
    assert 3.0e-5 < got["f16"][0] < 4.0e-5
    assert got["tf32"][0] < got["tf32_trunc"][0]  # rounding to nearest first is why the kernel does it
    # split-bf16 operands (16 mantissa bits, three MMAs per product): inside the fp32 variant's bars with room to spare
    y = orc.decode_features(GOLD["x"], WS, BS, operand="bf16x2")
    r = orc.relevancy(y, GOLD["phrases"], N_POS)[..., 0]
    e = (np.abs(y - GOLD["y"]).max(), np.abs(r - GOLD["relevancy"][..., 0]).max())
    assert e[0] < TOL["bf16x2"][0] / 4 and e[1] < TOL["bf16x2"][1] / 4, e


def test_host_validation_without_gpu():
    """FeatureDecoder refuses CPU tensors (no CPU path) and malformed layer lists before touching the library."""
    from langsplat_b200 import LsrError
    from langsplat_b200.decode import FeatureDecoder
    ws = [torch.from_numpy(w) for w in WS]
    bs = [torch.from_numpy(b) for b in BS]
    with pytest.raises(LsrError, match="CUDA tensors required"):
        FeatureDecoder(ws, bs)
    with pytest.raises(LsrError):
        FeatureDecoder(ws, bs[:-1])
    with pytest.raises(LsrError):
        FeatureDecoder([], [])
    with pytest.raises(ValueError):
        FeatureDecoder(ws, bs, precision="bf16")  # (the split variant is "bf16x2")
    with pytest.raises(LsrError, match="no decoder"):
        FeatureDecoder.from_state_dict({"encoder.0.weight": ws[0]}, device="cpu")


def test_abi_decoder_struct_matches_header():
    """ctypes mirror of lsr_decoder: field order / sizes as include/lsr_api.h declares them."""
    import ctypes
    from langsplat_b200 import _lib as L
    assert [f[0] for f in L.Decoder._fields_] == ["n_layers", "dims", "weight", "bias", "packed_tf32", "packed_f16",
                                                  "packed_bf16x2"]
    assert ctypes.sizeof(L.Decoder) == 4 + 4 * 9 + 8 * 8 + 8 * 8 + 8 + 8 + 8
    hdr = open(os.path.join(os.path.dirname(__file__), "..", "include", "lsr_api.h")).read()
    for sym in ("lsr_decoder_packed_bytes", "lsr_decoder_pack", "lsr_decode_work_bytes", "lsr_decode",
                "lsr_debug_decode_trace"):
        assert sym + "(" in hdr and sym in L.EXPORTS


def _decoder(precision):
    from langsplat_b200.decode import FeatureDecoder
    dev = torch.device("cuda:0")
    dec = FeatureDecoder([torch.from_numpy(w).to(dev) for w in WS], [torch.from_numpy(b).to(dev) for b in BS], precision)
    ph = torch.from_numpy(GOLD["phrases"]).to(dev)
    dec.set_phrases(ph[:N_POS], ph[N_POS:])
    return dec, dev


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["fp32", "bf16x2", "tf32", "f16"])
def test_gpu_matches_reference_vectors(precision):
    from langsplat_b200 import _lib
    dec, dev = _decoder(precision)
    x = torch.from_numpy(GOLD["x"]).to(dev)
    before = _lib.load().lsr_launch_count()
    y = dec.decode(x)
    rel = dec.relevancy_maps(x)
    torch.cuda.synchronize()
    assert _lib.load().lsr_launch_count() > before
    tf, tr = TOL[precision]
    assert np.abs(y.cpu().numpy() - GOLD["y"]).max() < tf
    assert np.abs(rel.cpu().numpy() - GOLD["relevancy"][..., 0]).max() < tr
    pair = dec.get_relevancy(x, 1).cpu().numpy()
    assert np.abs(pair - GOLD["relevancy"][1]).max() < tr


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["fp32", "bf16x2", "tf32", "f16"])
@pytest.mark.parametrize("n", [1, 31, 127, 128, 129, 1000, 148 * 128 * 2 + 77])
def test_gpu_matches_oracle_ragged(precision, n):
    dec, dev = _decoder(precision)
    rng = np.random.default_rng(100 + n)
    x = rng.normal(size=(n, 3)).astype(np.float32)
    x /= np.linalg.norm(x, axis=-1, keepdims=True)
    y0 = orc.decode_features(x, WS, BS)
    r0 = orc.relevancy(y0, GOLD["phrases"], N_POS)[..., 0]
    xt = torch.from_numpy(x).to(dev)
    tf, tr = TOL[precision]
    assert np.abs(dec.decode(xt).cpu().numpy() - y0).max() < tf
    assert np.abs(dec.relevancy_maps(xt).cpu().numpy() - r0).max() < tr


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["fp32", "bf16x2", "tf32", "f16"])
@pytest.mark.parametrize("n_pos,n_neg", [(1, 1), (12, 4), (2, 14)])
def test_gpu_phrase_counts(precision, n_pos, n_neg):
    """Fewest and most phrases the C ABI takes (LSR_DECODER_MAX_PHRASES = 16), against the oracle."""
    dec, dev = _decoder(precision)
    rng = np.random.default_rng(7 + n_pos)
    ph = rng.normal(size=(n_pos + n_neg, 512)).astype(np.float32)
    ph /= np.linalg.norm(ph, axis=-1, keepdims=True)
    x = rng.normal(size=(333, 3)).astype(np.float32)
    y0 = orc.decode_features(x, WS, BS)
    ph[0] = (0.5 * ph[0] + y0[:40].mean(0) * 10).astype(np.float32)
    ph[0] /= np.linalg.norm(ph[0])
    r0 = orc.relevancy(y0, ph, n_pos)[..., 0]
    pt = torch.from_numpy(ph).to(dev)
    dec.set_phrases(pt[:n_pos], pt[n_pos:])
    got = dec.relevancy_maps(torch.from_numpy(x).to(dev)).cpu().numpy()
    assert got.shape == (n_pos, 333)
    assert np.abs(got - r0).max() < TOL[precision][1]
    from langsplat_b200 import LsrError
    with pytest.raises(LsrError):
        dec.set_phrases(pt[:1].repeat(13, 1), pt[:1].repeat(4, 1))  # 17 phrases


@pytest.mark.gpu
def test_gpu_empty_and_errors():
    from langsplat_b200.decode import FeatureDecoder
    from langsplat_b200 import LsrError
    dec, dev = _decoder("tf32")
    assert dec.decode(torch.empty((0, 3), device=dev)).shape == (0, 512)
    assert dec.relevancy_maps(torch.empty((0, 3), device=dev)).shape == (N_POS, 0)
    with pytest.raises(LsrError):
        dec.decode(torch.zeros((4, 5), device=dev))
    with pytest.raises(LsrError):
        dec.decode(torch.zeros((4, 3)))
    odd = FeatureDecoder([torch.randn(10, 3, device=dev), torch.randn(7, 10, device=dev)],
                         [torch.randn(10, device=dev), torch.randn(7, device=dev)], "tf32")
    x = torch.randn(50, 3, device=dev)
    with pytest.raises(LsrError, match="multiples of 16"):
        odd.decode(x)
    y = odd.decode(x, precision="fp32")  # the fp32 variant takes any widths
    w = [t.cpu().numpy() for t in odd.weights]
    b = [t.cpu().numpy() for t in odd.biases]
    assert np.abs(y.cpu().numpy() - orc.decode_features(x.cpu().numpy(), w, b)).max() < 1e-5
    with pytest.raises(LsrError):
        odd.relevancy_maps(x)  # no phrases set


@pytest.mark.gpu
def test_gpu_full_size_properties():
    """A 1440x1080 eval image, three levels (the shape of eval/evaluate_iou_loc.py:256-259): unit outputs, both
    variants agree, get_max_across has the reference's layout, many tiles per CTA."""
    dec, dev = _decoder("tf32")
    g = torch.Generator(device="cpu").manual_seed(5)
    sem = torch.randn(3, 270, 360, 3, generator=g)
    sem = (sem / sem.norm(dim=-1, keepdim=True)).to(dev)
    y = dec.decode(sem)
    assert y.shape == (3, 270, 360, 512)
    assert float((y.norm(dim=-1) - 1).abs().max()) < 1e-5
    y32 = dec.decode(sem, precision="fp32")
    rel32 = dec.get_max_across(sem, precision="fp32")
    big = torch.randn(1080 * 1440, 3, generator=g).to(dev)
    for prec in ("bf16x2", "tf32", "f16"):
        y = dec.decode(sem, precision=prec)
        assert float((y.norm(dim=-1) - 1).abs().max()) < 1e-5
        assert float((y - y32).abs().max()) < TOL[prec][0]
        rel = dec.get_max_across(sem, precision=prec)
        assert rel.shape == (3, N_POS, 270, 360)
        assert float((rel - rel32).abs().max()) < TOL[prec][1]
        assert float(rel.min()) >= 0.0 and float(rel.max()) <= 1.0
        r1 = dec.relevancy_maps(big, precision=prec)
        r2 = dec.relevancy_maps(big[: 1080 * 720], precision=prec)
        assert torch.equal(r1[:, : 1080 * 720], r2)  # pixels are independent: a prefix gives the same values
