"""Differential tests against the UNMODIFIED reference CUDA extension (baseline/_ref, built for sm_100a by the
recipe in DESIGN.md) on the same GPU.  This is what pins parity: the reference ships no tests or golden vectors.

The reference returns its three opaque byte buffers; their layout (SURVEY.md Appendix B,
RAST/cuda_rasterizer/rasterizer_impl.cu:155-194 with 128-byte `obtain` alignment) is parsed here to compare radii,
tile keys, the sorted point list and tile ranges bit for bit.
"""
import numpy as np
import pytest
import torch

from helpers import R, cuda_forward_state, load_reference, ref_forward_buffers, rel_err, run_module, syn

pytestmark = pytest.mark.gpu

REF = load_reference()
needs_ref = pytest.mark.skipif(REF is None, reason="baseline/_ref reference extension not built")


def _ref_forward_buffers(scene, cam, include_feature=True):
    return ref_forward_buffers(REF, scene, cam, include_feature)


def _np(t):
    return t.detach().cpu().numpy()


@needs_ref
@pytest.mark.parametrize("cfg,P,W,H", [(1, None, None, None), (2, 100_000, 960, 540), (3, 60_000, 720, 540)])
def test_forward_bit_exact_against_reference(cfg, P, W, H):
    scene, cams = syn.make_config(cfg, P=P, W=W, H=H)
    cam = cams[0]
    ref = _ref_forward_buffers(scene, cam)
    color, feat, radii, state = cuda_forward_state(scene, cam)
    arr = R.debug_arrays(state)
    assert state["num_rendered"] == ref["num_rendered"]
    vis = _np(ref["radii"]) > 0
    np.testing.assert_array_equal(_np(radii), _np(ref["radii"]))
    np.testing.assert_array_equal(_np(arr["tiles_touched"]).view(np.uint32), ref["tiles_touched"])
    np.testing.assert_array_equal(_np(arr["keys"]).view(np.uint64), ref["keys"])
    np.testing.assert_array_equal(_np(arr["point_list"]).view(np.uint32), ref["point_list"])
    np.testing.assert_array_equal(_np(arr["ranges"]).view(np.uint32), ref["ranges"])
    rec = _np(arr["rec"])
    np.testing.assert_array_equal(rec[vis, 0:2].view(np.uint32), ref["means2D"].reshape(-1, 2)[vis].view(np.uint32))
    np.testing.assert_array_equal(_np(arr["depths"])[vis].view(np.uint32), ref["depths"][vis].view(np.uint32))
    np.testing.assert_array_equal(rec[vis][:, [2, 3, 4]].view(np.uint32),
                                  ref["conic_opacity"].reshape(-1, 4)[vis][:, 0:3].view(np.uint32))
    np.testing.assert_array_equal(_np(arr["cov3D"])[vis].view(np.uint32), ref["cov3D"].reshape(-1, 6)[vis].view(np.uint32))
    np.testing.assert_array_equal(_np(arr["n_contrib"]).reshape(-1).view(np.uint32), ref["n_contrib"])
    assert np.abs(rec[vis][:, 6:9] - ref["rgb"].reshape(-1, 3)[vis]).max() < 2e-6
    assert (color - ref["color"]).abs().max().item() < 1e-5
    assert (feat - ref["feature"]).abs().max().item() < 1e-5
    assert np.abs(_np(arr["final_T"]).reshape(-1) - ref["final_T"]).max() < 1e-5


@needs_ref
@pytest.mark.parametrize("cfg,P,W,H,bg", [(1, None, None, None, (0, 0, 0)), (2, 100_000, 960, 540, (0.1, 0.4, 0.7))])
def test_gradients_against_reference(cfg, P, W, H, bg):
    scene, cams = syn.make_config(cfg, P=P, W=W, H=H)
    cam = cams[0]
    dcol, dfeat = syn.make_grads(cam.image_width, cam.image_height, 2000 + cfg)
    ours = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, grads=(dcol, dfeat), bg=bg)
    ref = run_module(REF.GaussianRasterizer, REF.GaussianRasterizationSettings, scene, cam, grads=(dcol, dfeat), bg=bg)
    assert torch.equal(ours["radii"], ref["radii"])
    assert (ours["color"] - ref["color"]).abs().max().item() < 1e-5
    assert (ours["feature"] - ref["feature"]).abs().max().item() < 1e-5
    for k in ("means3D", "means2D", "shs", "language_feature", "opacities", "scales", "rotations"):
        err = rel_err(ours["grads"][k], ref["grads"][k])
        assert err < 1e-4, f"{k}: rel err {err:.3e}"


@needs_ref
def test_feature_only_training_against_reference():
    scene, cams = syn.make_config(3, P=60_000, W=720, H=540)
    cam = cams[0]
    _, dfeat = syn.make_grads(cam.image_width, cam.image_height, 2003, feature_only=True)
    kw = dict(grads=(None, dfeat), requires=("language_feature",))
    ours = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, **kw)
    ref = run_module(REF.GaussianRasterizer, REF.GaussianRasterizationSettings, scene, cam, **kw)
    assert (ours["feature"] - ref["feature"]).abs().max().item() < 1e-5
    for k in ("language_feature", "means2D"):
        err = rel_err(ours["grads"][k], ref["grads"][k])
        assert err < 1e-4, f"{k}: rel err {err:.3e}"
