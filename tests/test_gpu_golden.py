"""GPU: the CUDA path (public module -> C ABI -> liblsr.so) against the reference extension's golden vectors
(tests/golden/*.npz).  Needs no reference build and no oracle: the fixtures carry the reference's own outputs."""
import numpy as np
import pytest
import torch

from helpers import GOLDEN_CASES, R, cuda_forward_state, golden_available, load_golden, rel_err, run_module

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not golden_available(), reason="tests/golden/*.npz missing")]


def _np(t):
    return t.detach().cpu().numpy()


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_cuda_forward_matches_golden(name):
    fx, scene, cam, kw = load_golden(name)
    color, feat, radii, state = cuda_forward_state(scene, cam, **kw)
    arr = R.debug_arrays(state)
    vis = fx["radii"] > 0
    assert state["num_rendered"] == int(fx["num_rendered"])
    np.testing.assert_array_equal(_np(radii), fx["radii"])
    np.testing.assert_array_equal(_np(arr["tiles_touched"]).view(np.uint32), fx["tiles_touched"])
    np.testing.assert_array_equal(_np(arr["keys"]).view(np.uint64), fx["keys"])
    np.testing.assert_array_equal(_np(arr["point_list"]).view(np.uint32), fx["point_list"])
    np.testing.assert_array_equal(_np(arr["ranges"]).view(np.uint32), fx["ranges"])
    np.testing.assert_array_equal(_np(arr["n_contrib"]).reshape(-1).view(np.uint32), fx["n_contrib"])
    rec = _np(arr["rec"])
    np.testing.assert_array_equal(rec[vis, 0:2].view(np.uint32), fx["means2D"].reshape(-1, 2)[vis].view(np.uint32))
    np.testing.assert_array_equal(_np(arr["depths"])[vis].view(np.uint32), fx["depths"][vis].view(np.uint32))
    np.testing.assert_array_equal(rec[vis][:, [2, 3, 4]].view(np.uint32),
                                  fx["conic_opacity"].reshape(-1, 4)[vis][:, 0:3].view(np.uint32))
    np.testing.assert_allclose(_np(color), fx["color"], atol=1e-5, rtol=0)
    if kw["include_feature"]:
        np.testing.assert_allclose(_np(feat), fx["feature"], atol=1e-5, rtol=0)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_cuda_backward_matches_golden(name):
    fx, scene, cam, kw = load_golden(name)
    feature_only = bool(fx["in_feature_only"])
    requires = tuple(k[5:] for k in fx if k.startswith("grad_") and k != "grad_means2D")
    dcol = None if feature_only else torch.from_numpy(fx["in_dL_dcolor"])
    dfeat = torch.from_numpy(fx["in_dL_dfeature"]) if kw["include_feature"] else None
    out = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, grads=(dcol, dfeat),
                     requires=requires, **kw)
    for k in fx:
        if not k.startswith("grad_"):
            continue
        ours = out["grads"][k[5:]]
        assert ours is not None, k
        err = rel_err(ours, torch.from_numpy(fx[k]))
        assert err < 1e-4, f"{name}/{k}: rel err {err:.3e}"
