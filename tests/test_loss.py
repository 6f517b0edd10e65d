"""Fused L1 loss: the oracle against golden vectors produced by the reference's own `l1_loss`
(tests/golden/make_golden_loss.py), and — on the GPU — the CUDA kernels (public `l1_loss` -> C ABI) against the
oracle, the golden vectors and PyTorch's own op sequence at full image size.
Tolerances: the loss is a mean of |.| terms summed in a different order -> 1e-6 relative; the gradient is an exact
product of the sign, the mask and a scalar -> bit-exact up to the rounding of grad_out/numel (1 ulp)."""
import os

import numpy as np
import pytest
import torch

from oracle import oracle as orc

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "l1_loss.npz")
CASES = ["rgb_33x17", "feat_masked_40x24", "feat_masked_odd_19x7", "ties_16x16"]


def _case(name):
    z = np.load(GOLD)
    mask = z[name + ".mask"] if name + ".mask" in z.files else None
    return z[name + ".x"], z[name + ".gt"], mask, float(z[name + ".gscale"]), z[name + ".loss"], z[name + ".grad"]


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_l1_loss(name):
    x, gt, mask, gs, loss, grad = _case(name)
    assert abs(float(orc.l1_loss(x, gt, mask)) - float(loss)) <= 1e-6 * abs(float(loss))
    np.testing.assert_allclose(orc.l1_loss_grad(x, gt, mask, gs), grad, rtol=2e-7, atol=0)


def test_oracle_l1_edge_cases():
    x = np.zeros((3, 0, 5), np.float32)
    assert np.isnan(orc.l1_loss(x, x))
    x = np.ones((3, 4, 4), np.float32)
    assert orc.l1_loss(x, x) == 0 and not orc.l1_loss_grad(x, x).any()  # sign(0) = 0
    m = np.zeros((1, 4, 4), np.float32)
    assert orc.l1_loss(x, 2 * x, m) == 0 and not orc.l1_loss_grad(x, 2 * x, m).any()


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_l1_loss_matches_golden_and_oracle(name):
    from langsplat_b200 import l1_loss
    x, gt, mask, gs, loss, grad = _case(name)
    dev = torch.device("cuda:0")
    xt = torch.from_numpy(x).to(dev).requires_grad_(True)
    mt = None if mask is None else torch.from_numpy(mask).to(dev)
    out = l1_loss(xt, torch.from_numpy(gt).to(dev), mt)
    (out * gs).backward()
    assert abs(out.item() - float(loss)) <= 1e-6 * abs(float(loss))
    assert abs(out.item() - float(orc.l1_loss(x, gt, mask))) <= 1e-6 * abs(float(loss))
    np.testing.assert_allclose(xt.grad.cpu().numpy(), grad, rtol=2e-7, atol=0)
    np.testing.assert_allclose(xt.grad.cpu().numpy(), orc.l1_loss_grad(x, gt, mask, gs), rtol=2e-7, atol=0)


@pytest.mark.gpu
@pytest.mark.parametrize("shape,masked", [((3, 1080, 1920), True), ((3, 1080, 1920), False), ((3, 731, 1237), True),
                                          ((1, 1, 1), False)])
def test_cuda_l1_loss_full_size_vs_torch(shape, masked):
    """BASELINE-size images: the fused kernels against PyTorch's own op sequence (train.py:96-99) on the same GPU."""
    from langsplat_b200 import l1_loss
    dev = torch.device("cuda:0")
    g = torch.Generator(device="cpu").manual_seed(5)
    x = torch.rand(shape, generator=g).to(dev).requires_grad_(True)
    gt = torch.rand(shape, generator=g).to(dev)
    mask = (torch.rand((1,) + shape[1:], generator=g) > 0.25).float().to(dev) if masked else None
    ours = l1_loss(x, gt, mask)
    ours.backward()
    g_ours = x.grad.clone()
    x.grad = None
    ref = torch.abs(x * mask - gt * mask).mean() if masked else torch.abs(x - gt).mean()
    ref.backward()
    assert abs(ours.item() - ref.item()) <= 2e-6 * abs(ref.item())
    np.testing.assert_allclose(g_ours.cpu().numpy(), x.grad.cpu().numpy(), rtol=2e-7, atol=0)


@pytest.mark.gpu
def test_cuda_l1_loss_errors_and_empty():
    from langsplat_b200 import LsrError, l1_loss
    dev = torch.device("cuda:0")
    with pytest.raises(LsrError):
        l1_loss(torch.zeros(3, 4, 4), torch.zeros(3, 4, 4))  # CPU tensors: no CPU path
    with pytest.raises(LsrError):
        l1_loss(torch.zeros(3, 4, 4, device=dev), torch.zeros(3, 4, 5, device=dev))
    with pytest.raises(LsrError):
        l1_loss(torch.zeros(3, 4, 4, device=dev), torch.zeros(3, 4, 4, device=dev), torch.zeros(5, device=dev))
    assert torch.isnan(l1_loss(torch.zeros(3, 0, 4, device=dev), torch.zeros(3, 0, 4, device=dev)))
