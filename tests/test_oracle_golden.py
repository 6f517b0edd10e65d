"""CPU: the oracle (oracle/oracle.cpp) against the golden vectors the UNMODIFIED reference CUDA extension produced on a
B200 (tests/golden/*.npz, generator: tests/golden/make_golden.py).  This is what pins the oracle: the reference ships
no tests or fixtures of its own.

Bars: bit-exact radii, depths, means2D, cov3D, conic, tiles_touched, keys, sorted point list, tile ranges, n_contrib;
images and final_T within 1e-5 max-abs; rgb within 2e-6 (glibc vs CUDA expf/sqrt ulps are not on that chain, SH is a
plain polynomial); gradients within 1e-4 relative (max-abs error / max-abs value)."""
import numpy as np
import pytest
import torch

from helpers import GOLDEN_CASES, golden_available, load_golden, oracle_forward, orc, rel_err

pytestmark = pytest.mark.skipif(not golden_available(), reason="tests/golden/*.npz not generated yet")

IMG_TOL, GRAD_TOL = 1e-5, 1e-4


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_oracle_forward_matches_reference_golden(name):
    fx, scene, cam, kw = load_golden(name)
    fw = oracle_forward(scene, cam, **kw)
    vis = fx["radii"] > 0
    assert fw["num_rendered"] == int(fx["num_rendered"])
    np.testing.assert_array_equal(fw["radii"], fx["radii"])
    np.testing.assert_array_equal(fw["tiles_touched"], fx["tiles_touched"])
    np.testing.assert_array_equal(fw["depths"][vis].view(np.uint32), fx["depths"][vis].view(np.uint32))
    np.testing.assert_array_equal(fw["means2D"][vis].view(np.uint32), fx["means2D"].reshape(-1, 2)[vis].view(np.uint32))
    np.testing.assert_array_equal(fw["cov3D"][vis].view(np.uint32), fx["cov3D"].reshape(-1, 6)[vis].view(np.uint32))
    np.testing.assert_array_equal(fw["conic_opacity"][vis].view(np.uint32),
                                  fx["conic_opacity"].reshape(-1, 4)[vis].view(np.uint32))
    np.testing.assert_array_equal(fw["keys"], fx["keys"])
    np.testing.assert_array_equal(fw["point_list"], fx["point_list"])
    np.testing.assert_array_equal(fw["ranges"], fx["ranges"])
    np.testing.assert_array_equal(fw["n_contrib"].reshape(-1), fx["n_contrib"])
    np.testing.assert_allclose(fw["rgb"][vis], fx["rgb"].reshape(-1, 3)[vis], atol=2e-6, rtol=0)
    np.testing.assert_allclose(fw["final_T"].reshape(-1), fx["final_T"], atol=IMG_TOL, rtol=0)
    np.testing.assert_allclose(fw["color"], fx["color"], atol=IMG_TOL, rtol=0)
    if kw["include_feature"]:
        np.testing.assert_allclose(fw["feature"], fx["feature"], atol=IMG_TOL, rtol=0)
    else:
        assert fx["feature"].shape == (1,) and fx["feature"][0] == 0  # rasterize_points.cu:72-77


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_oracle_backward_matches_reference_golden(name):
    fx, scene, cam, kw = load_golden(name)
    fw = oracle_forward(scene, cam, **kw)
    feature_only = bool(fx["in_feature_only"])
    dcol = np.zeros_like(fx["in_dL_dcolor"]) if feature_only else fx["in_dL_dcolor"]
    g = orc.backward(fw, dcol, fx["in_dL_dfeature"] if kw["include_feature"] else None)
    names = {"means3D": "means3D", "means2D": "means2D", "shs": "shs", "language_feature": "language_feature",
             "opacities": "opacities", "scales": "scales", "rotations": "rotations"}
    checked = 0
    for k_ref, k_or in names.items():
        key = "grad_" + k_ref
        if key not in fx:
            continue
        ref = torch.from_numpy(fx[key])
        ours = torch.from_numpy(np.asarray(g[k_or])).reshape(ref.shape)
        err = rel_err(ours, ref)
        assert err < GRAD_TOL, f"{name}/{k_ref}: rel err {err:.3e}"
        checked += 1
    assert checked >= 2
