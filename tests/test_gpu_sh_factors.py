"""GPU: factored SH gradients (langsplat_b200.dist.factored_sh_gradients / finish_sh_gradients ->
lsr_grads.dL_dcolors_masked + lsr_sh_gradient_from_views).  One view's dL_dsh is basis(dir) (x) g per Gaussian
(RAST cuda_rasterizer/backward.cu:31-129 restated in oracle/oracle.py), so the sum over the views of a step formed from
the 12-byte factors must equal the sum of the dense per-view gradients up to float summation order, every other
gradient must be unchanged by the factoring, and the oracle's SH gradient pins both."""
import numpy as np
import pytest
import torch

from helpers import R, oracle_forward, orc, rel_err, settings_for, syn
from langsplat_b200 import dist as D

pytestmark = pytest.mark.gpu

NAMES = ("means3D", "shs", "opacities", "scales", "rotations", "language_feature")


def _step(scene, cams, dev, degree, factored):
    params = {k: scene[k].to(dev).requires_grad_(True) for k in NAMES}
    means2D = torch.zeros_like(params["means3D"], requires_grad=True)

    def run():
        for i, cam in enumerate(cams):
            dcol, dfeat = [t.to(dev) for t in syn.make_grads(cam.image_width, cam.image_height, 300 + i)]
            rast = R.GaussianRasterizer(settings_for(cam, dev, sh_degree=degree))
            color, feat, radii = rast(means3D=params["means3D"], means2D=means2D, shs=params["shs"],
                                      language_feature_precomp=params["language_feature"],
                                      opacities=params["opacities"], scales=params["scales"],
                                      rotations=params["rotations"])
            torch.autograd.backward([color, feat], [dcol, dfeat])

    if factored:
        with D.factored_sh_gradients() as factors:
            run()
        assert params["shs"].grad is None and len(factors) == len(cams)
        params["shs"].grad = D.finish_sh_gradients(factors, params["means3D"], params["shs"].shape[1])
    else:
        run()
    return {k: p.grad.detach().cpu() for k, p in params.items()}


@pytest.mark.parametrize("degree", [0, 1, 2, 3])
def test_factored_step_equals_dense_step(degree):
    dev = torch.device("cuda:0")
    scene, cams = syn.make_config(4, P=20000, W=200, H=136, n_cameras=3)
    dense = _step(scene, cams, dev, degree, factored=False)
    again = _step(scene, cams, dev, degree, factored=False)
    fact = _step(scene, cams, dev, degree, factored=True)
    # same kernels, same inputs; the blend backward sums with unordered atomics, so two dense runs already differ in
    # the last bits: the bar is that run-to-run noise (floored at 1e-5), far below what a missing term would show
    for k in NAMES:
        if k != "shs":
            tol = max(1e-5, 4.0 * rel_err(again[k], dense[k]))
            assert rel_err(fact[k], dense[k]) <= tol, f"{k} changed under SH factoring"
    assert fact["shs"].shape == dense["shs"].shape
    n = (degree + 1) ** 2
    assert float(fact["shs"][:, n:].abs().max()) == 0.0 if n < 16 else True  # unused bands: exact zeros, like dense
    assert rel_err(fact["shs"], dense["shs"]) <= max(1e-5, 4.0 * rel_err(again["shs"], dense["shs"]))


def test_single_view_factor_matches_oracle():
    dev = torch.device("cuda:0")
    scene, cams = syn.make_config(1, P=6000, W=128, H=96)
    cam = cams[0]
    got = _step(scene, [cam], dev, 3, factored=True)
    dcol, dfeat = syn.make_grads(cam.image_width, cam.image_height, 300)
    fw = oracle_forward(scene, cam)
    g = orc.backward(fw, dcol.numpy(), dfeat.numpy())
    want = np.asarray(g["shs"], np.float32).reshape(scene["shs"].shape)
    assert rel_err(got["shs"], want) <= 1e-4  # the gradient bar of the path


def test_no_views_and_bad_arguments():
    dev = torch.device("cuda:0")
    means = torch.randn(100, 3, device=dev)
    assert D.finish_sh_gradients([], means, 16) is None
    from langsplat_b200 import _lib as L
    import ctypes
    lib = L.load()
    out = torch.empty(100, 4, 3, device=dev)
    rc = lib.lsr_sh_gradient_from_views(100, 4, 3, 0, ctypes.c_void_p(means.data_ptr()), None, None,
                                        ctypes.c_void_p(out.data_ptr()), None)
    assert rc != L.LSR_OK and b"SH coefficients" in L.last_error().encode()
    out = torch.full((100, 16, 3), 7.0, device=dev)  # zero views: the sum is zero
    rc = lib.lsr_sh_gradient_from_views(100, 16, 3, 0, ctypes.c_void_p(means.data_ptr()), None, None,
                                        ctypes.c_void_p(out.data_ptr()),
                                        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == L.LSR_OK
    torch.cuda.synchronize()
    assert float(out.abs().max()) == 0.0
