"""Fused L1 loss: the oracle against golden vectors produced by the reference's own `l1_loss`
(tests/golden/make_golden_loss.py), and — on the GPU — the CUDA kernels (public `l1_loss` -> C ABI) against the
oracle, the golden vectors and PyTorch's own op sequence at full image size.
Tolerances: the loss is a mean of |.| terms summed in a different order -> 1e-6 relative on the small cases, 4e-6 at
full image size (the per-CTA partial sums are added by float atomics in completion order); the gradient is an exact
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
    assert abs(ours.item() - ref.item()) <= 4e-6 * abs(ref.item())  # 6.2 M terms, per-CTA partials added by float atomics
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


@pytest.mark.gpu
def test_cuda_l1_loss_takes_the_bool_mask_train_py_passes():
    """scene/cameras.py:70 returns `seg != -1` (bool) and train.py:96-99 multiplies it in as it is; so does this
    repo's drop-in get_language_feature.  The mask goes straight into l1_loss."""
    from langsplat_b200 import l1_loss
    from langsplat_b200.gt_features import language_feature_map
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(11)
    H, W, K = 37, 52, 9
    seg = torch.randint(-1, K, (4, H, W), generator=g).float().to(dev)
    table = torch.randn(K, 3, generator=g).to(dev)
    gt, mask = language_feature_map(seg, table, 1, H, W)
    assert mask.dtype == torch.bool
    x = torch.rand(3, H, W, generator=g).to(dev).requires_grad_(True)
    ours = l1_loss(x, gt, mask)
    ours.backward()
    g_ours = x.grad.clone()
    x.grad = None
    ref = torch.abs(x * mask - gt * mask).mean()
    ref.backward()
    assert abs(ours.item() - ref.item()) <= 4e-6 * abs(ref.item())  # 6.2 M terms, per-CTA partials added by float atomics
    np.testing.assert_allclose(g_ours.cpu().numpy(), x.grad.cpu().numpy(), rtol=2e-7, atol=0)
    assert abs(l1_loss(x, gt, mask.to(torch.uint8)).item() - ref.item()) <= 2e-6 * abs(ref.item())


@pytest.mark.gpu
def test_cuda_l1_loss_misaligned_views():
    """Contiguous views that do not start on a 16-byte boundary (storage offset 1) must take the scalar path, not
    fault: the float4 fast path checks pointer alignment, not only the shape."""
    from langsplat_b200 import l1_loss
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(12)
    n = 3 * 16 * 16
    base_x, base_gt = torch.rand(n + 1, generator=g).to(dev), torch.rand(n + 1, generator=g).to(dev)
    base_m = (torch.rand(16 * 16 + 1, generator=g) > 0.3).float().to(dev)
    x = base_x[1:].view(3, 16, 16).requires_grad_(True)
    gt, m = base_gt[1:].view(3, 16, 16), base_m[1:].view(1, 16, 16)
    assert x.data_ptr() % 16 == 4 and x.is_contiguous()
    ours = l1_loss(x, gt, m)
    ours.backward()
    ref = torch.abs(x.detach() * m - gt * m).mean()
    assert abs(ours.item() - ref.item()) <= 2e-6 * abs(ref.item())
    want = torch.sign(x.detach() * m - gt * m) * m / n
    np.testing.assert_allclose(x.grad.cpu().numpy(), want.cpu().numpy(), rtol=2e-7, atol=0)


@pytest.mark.gpu
@pytest.mark.parametrize("shape,masked,offset", [((3, 1080, 1920), False, 0), ((3, 540, 960), True, 0),
                                                 ((3, 37, 53), False, 0), ((3, 16, 16), True, 1)])
def test_cuda_l1_loss_uint8_ground_truth(shape, masked, offset):
    """gt handed over as the stored 8-bit image: the kernels form u / 255 themselves (IEEE division, the float value
    utils/general_utils.py:23 computes on the host).  Must equal the float path on `gt_u8 / 255.0` bit for bit in the
    gradient, and to summation order in the loss; offset 1 = a view that is not 4-byte aligned (scalar path)."""
    from langsplat_b200 import l1_loss
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(21)
    n = shape[0] * shape[1] * shape[2]
    x = torch.rand(shape, generator=g).to(dev).requires_grad_(True)
    raw = torch.randint(0, 256, (n + offset,), generator=g, dtype=torch.uint8).to(dev)
    gt_u8 = raw[offset:].view(shape)
    mask = (torch.rand((1,) + shape[1:], generator=g) > 0.25).to(dev) if masked else None
    gt_f = (gt_u8.cpu() / 255.0).to(dev)  # PILtoTorch's arithmetic
    assert gt_f.dtype == torch.float32
    a = l1_loss(x, gt_u8, mask)
    a.backward()
    ga = x.grad.clone()
    x.grad = None
    b = l1_loss(x, gt_f, mask)
    b.backward()
    # same per-element values; the per-CTA partial sums meet in one float atomicAdd in whatever order the CTAs finish
    # (24 k of them at 1080p): two runs of the SAME kernel already differ by ~1e-6 relative
    assert abs(a.item() - b.item()) <= 4e-6 * abs(b.item())
    assert torch.equal(ga, x.grad)
    m = mask.float() if masked else torch.ones((), device=dev)
    ref = torch.abs(x.detach() * m - gt_f * m).mean()
    assert abs(a.item() - ref.item()) <= 4e-6 * abs(ref.item())
    with pytest.raises(Exception):
        l1_loss(x, gt_u8.cpu(), mask)
