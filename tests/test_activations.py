"""Fused activations in front of the rasterizer: the oracle's restatement against golden vectors
produced by the reference's own GaussianModel properties (tests/golden/make_golden_activations.py), then — on the GPU —
lsr_activate_forward / backward through langsplat_b200.activate against the same vectors and against the oracle on
larger seeded inputs.  Tolerances: exp and sigmoid are the same float expressions (1 ulp of slack for libm vs CUDA
expf); the two normalisations sum their squares in a fixed left-to-right order that may differ from PyTorch's
reduction in the last bit: 2e-7 relative to the row's largest output.  Gradients: 1e-5 relative (max-abs / max-abs)."""
import os

import numpy as np
import pytest
import torch

from helpers import orc, rel_err

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "activations.npz"))
CASES = ("mixed_257", "edge_64")
KEYS = ("scaling", "rotation", "opacity", "feature")


def _case(name):
    return {k: {f: GOLD[f"{name}.{k}.{f}"] for f in ("raw", "out", "up", "grad")} for k in KEYS}


def _close(got, want, tol):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    assert got.shape == want.shape
    scale = np.maximum(np.abs(want), np.finfo(np.float32).tiny)
    assert np.all(np.isfinite(got) == np.isfinite(want))
    ok = np.isfinite(want)
    assert float((np.abs(got - want)[ok] / scale[ok]).max(initial=0.0)) <= tol


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_vectors(name):
    c = _case(name)
    outs = orc.activate(*[c[k]["raw"] for k in KEYS])
    for k, y in zip(KEYS, outs):
        _close(y, c[k]["out"], 3e-7)
    grads = orc.activate_grad(*[c[k]["raw"] for k in KEYS], *[c[k]["up"] for k in KEYS])
    for k, g in zip(KEYS, grads):
        assert rel_err(g, c[k]["grad"]) <= 1e-6, k


def test_oracle_partial_arguments():
    c = _case("mixed_257")
    s, r, o, f = orc.activate(opacity=c["opacity"]["raw"])
    assert s is None and r is None and f is None and o.shape == (257, 1)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_gpu_matches_reference_vectors(name):
    from langsplat_b200 import activate
    dev = torch.device("cuda:0")
    c = _case(name)
    raw = [torch.from_numpy(c[k]["raw"]).to(dev).requires_grad_(True) for k in KEYS]
    outs = activate(*raw)
    for k, y in zip(KEYS, outs):
        _close(y.detach().cpu().numpy(), c[k]["out"], 3e-7)
    torch.autograd.backward(list(outs), [torch.from_numpy(c[k]["up"]).to(dev) for k in KEYS])
    for k, t in zip(KEYS, raw):
        assert rel_err(t.grad, c[k]["grad"]) <= 1e-5, k
        assert torch.isfinite(t.grad).all()


@pytest.mark.gpu
def test_gpu_matches_oracle_large_and_partial():
    from langsplat_b200 import LsrError, activate
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(11)
    P = 100003
    raw = [torch.randn(P, 3, generator=g) - 2.0, torch.randn(P, 4, generator=g), 3.0 * torch.randn(P, 1, generator=g),
           torch.randn(P, 3, generator=g)]
    ups = [torch.randn(t.shape, generator=g) for t in raw]
    want = orc.activate(*[t.numpy() for t in raw])
    wgrad = orc.activate_grad(*[t.numpy() for t in raw], *[u.numpy() for u in ups])
    dev_raw = [t.to(dev).requires_grad_(True) for t in raw]
    outs = activate(*dev_raw)
    for y, w in zip(outs, want):
        _close(y.detach().cpu().numpy(), w, 3e-7)
    torch.autograd.backward(list(outs), [u.to(dev) for u in ups])
    for t, w in zip(dev_raw, wgrad):
        assert rel_err(t.grad, w) <= 1e-5
    # the feature-only training step activates everything but differentiates the language feature alone
    lang = raw[3].to(dev).requires_grad_(True)
    s, r, o, f = activate(raw[0].to(dev), raw[1].to(dev), raw[2].to(dev), lang)
    assert torch.equal(s, outs[0].detach()) and torch.equal(f, outs[3].detach())
    f.backward(ups[3].to(dev))
    assert rel_err(lang.grad, wgrad[3]) <= 1e-5
    none_s, none_r, o2, none_f = activate(opacity=raw[2].to(dev))
    assert none_s is None and none_r is None and none_f is None and torch.equal(o2, outs[2].detach())
    with pytest.raises(LsrError):
        activate(raw[0].to(dev), raw[1][:5].to(dev))
    with pytest.raises(LsrError):
        activate(raw[0])  # CPU tensor: no CPU path
    with pytest.raises(LsrError):
        activate()
