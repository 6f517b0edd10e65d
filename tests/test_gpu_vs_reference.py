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


@needs_ref
@pytest.mark.parametrize("cfg,P,W,H,bg", [(1, None, None, None, (0, 0, 0)), (2, 200_000, 960, 540, (0.1, 0.4, 0.7)),
                                          (3, 100_000, 720, 540, (1, 1, 1))])
def test_images_are_bit_identical_to_the_reference(cfg, P, W, H, bg):
    """Stronger than the north star's 1e-5: with the default build the forward keeps the reference's rounding order
    all the way (SH colour, conic, power, exp, alpha, the per-channel (c * alpha) * T blend), so the images come out
    bit for bit — also measured at full C2 / C3 size (profiles/r2ae/image_diff_main.log: 0 of 6.2 M pixels differ)."""
    import os
    if os.environ.get("LSR_FAST_BLEND", "0")[:1] == "1":
        pytest.skip("LSR_FAST_BLEND=1 trades the bit-identical images for speed")
    scene, cams = syn.make_config(cfg, P=P, W=W, H=H)
    ours = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cams[0], bg=bg)
    ref = run_module(REF.GaussianRasterizer, REF.GaussianRasterizationSettings, scene, cams[0], bg=bg)
    assert torch.equal(ours["radii"], ref["radii"])
    assert torch.equal(ours["color"], ref["color"]), (ours["color"] - ref["color"]).abs().max().item()
    assert torch.equal(ours["feature"], ref["feature"]), (ours["feature"] - ref["feature"]).abs().max().item()


@needs_ref
def test_fast_blend_opt_in_stays_inside_the_image_tolerance():
    """LSR_FAST_BLEND=1 (read once per process, hence the subprocess): one blend weight alpha * T per pair instead of
    the reference's per-channel order.  Images within 1e-5 of the reference (measured 2.4e-7), everything else — radii,
    n_contrib, final T, gradients' bar — as before."""
    import os
    import subprocess
    import sys
    code = """
import sys, torch
sys.path.insert(0, %r)
from helpers import R, load_reference, run_module, syn, rel_err
REF = load_reference()
scene, cams = syn.make_config(2, P=200_000, W=960, H=540)
dcol, dfeat = syn.make_grads(cams[0].image_width, cams[0].image_height, 31)
ours = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cams[0], grads=(dcol, dfeat))
ref = run_module(REF.GaussianRasterizer, REF.GaussianRasterizationSettings, scene, cams[0], grads=(dcol, dfeat))
assert torch.equal(ours["radii"], ref["radii"])
dc = (ours["color"] - ref["color"]).abs().max().item()
df = (ours["feature"] - ref["feature"]).abs().max().item()
g = max(rel_err(ours["grads"][k], ref["grads"][k]) for k in ("means3D", "shs", "language_feature", "opacities", "scales", "rotations"))
print("RESULT", dc, df, g)
""" % os.path.dirname(os.path.abspath(__file__))
    env = dict(os.environ, LSR_FAST_BLEND="1")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    dc, df, g = [float(v) for v in out.stdout.strip().splitlines()[-1].split()[1:]]
    assert 0.0 < dc < 1e-5 and 0.0 < df < 1e-5, (dc, df)  # > 0: the switch took effect
    assert g < 1e-4
