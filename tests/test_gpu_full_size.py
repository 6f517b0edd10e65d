"""GPU, BASELINE.json's full sizes: the CUDA path against the live reference extension where it is built, plus
size-independent properties (sortedness, ranges partition the list, n_contrib bounded, capacity-retry idempotence)."""
import numpy as np
import pytest
import torch

from helpers import (R, compare_with_reference_runs, cuda_forward_state, load_reference, oracle_step_f64,
                     ref_forward_buffers, rel_err, run_module, syn)

pytestmark = pytest.mark.gpu

REF = load_reference()


def _np(t):
    return t.detach().cpu().numpy()


def _check_structure(arr, state, W, H):
    Rn = state["num_rendered"]
    keys = _np(arr["keys"]).view(np.uint64)
    assert keys.size == Rn
    # sorted by (tile, depth bits); the list only names visible Gaussians
    assert np.all(keys[1:] >= keys[:-1])
    tiles = (keys >> np.uint64(32)).astype(np.int64)
    ranges = _np(arr["ranges"]).astype(np.int64)
    gx, gy = (W + 15) // 16, (H + 15) // 16
    T = gx * gy
    assert tiles.max() < T
    lens = ranges[:, 1] - ranges[:, 0]
    assert lens.min() >= 0 and lens.sum() == Rn
    counts = np.bincount(tiles, minlength=T)
    np.testing.assert_array_equal(lens, counts)
    nz = lens > 0
    np.testing.assert_array_equal(ranges[nz, 0], (np.cumsum(counts) - counts)[nz])
    # the key's depth word is the depth of the listed Gaussian
    plist = _np(arr["point_list"]).view(np.uint32)
    depth_bits = _np(arr["depths"]).view(np.uint32)
    np.testing.assert_array_equal((keys & np.uint64(0xFFFFFFFF)).astype(np.uint32), depth_bits[plist])
    assert (_np(state["radii"]) > 0)[plist].all()
    # n_contrib never exceeds the tile's list length
    nc = _np(arr["n_contrib"]).astype(np.int64)
    per_pixel_len = np.kron(lens.reshape(gy, gx), np.ones((16, 16), np.int64))
    assert np.all(nc <= per_pixel_len[:H, :W])


def test_config2_full_size_structure_and_reference():
    scene, cams = syn.make_config(2)
    cam = cams[0]
    color, feat, radii, state = cuda_forward_state(scene, cam)
    arr = R.debug_arrays(state)
    assert torch.isfinite(color).all() and torch.isfinite(feat).all()
    _check_structure(arr, state, cam.image_width, cam.image_height)
    if REF is None:
        pytest.skip("baseline/_ref not built: structure checked only")
    ref = ref_forward_buffers(REF, scene, cam)
    assert state["num_rendered"] == ref["num_rendered"]
    np.testing.assert_array_equal(_np(radii), _np(ref["radii"]))
    np.testing.assert_array_equal(_np(arr["keys"]).view(np.uint64), ref["keys"])
    np.testing.assert_array_equal(_np(arr["point_list"]).view(np.uint32), ref["point_list"])
    np.testing.assert_array_equal(_np(arr["ranges"]).view(np.uint32), ref["ranges"])
    np.testing.assert_array_equal(_np(arr["n_contrib"]).reshape(-1).view(np.uint32), ref["n_contrib"])
    assert (color - ref["color"]).abs().max().item() < 1e-5
    assert (feat - ref["feature"]).abs().max().item() < 1e-5


GRAD_KEYS = ("means3D", "means2D", "shs", "language_feature", "opacities", "scales", "rotations")


@pytest.mark.skipif(REF is None, reason="baseline/_ref reference extension not built")
def test_config2_full_size_gradients_against_reference():
    scene, cams = syn.make_config(2)
    cam = cams[0]
    dcol, dfeat = syn.make_grads(cam.image_width, cam.image_height, 2002)
    ours = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, grads=(dcol, dfeat))
    refs = [run_module(REF.GaussianRasterizer, REF.GaussianRasterizationSettings, scene, cam, grads=(dcol, dfeat))
            for _ in range(3)]
    # The bar is the north star's 1e-4 (max-abs error / max-abs value).  At this size the reference's unordered float
    # atomics make its OWN runs differ by 3e-5 ... 1.2e-4 on dL_dscales (profiles/tools/grad_noise.py; every other
    # tensor stays below 5e-6; ours vs ours: < 2e-5), so a single reference run is not a fixed point to measure
    # against: compare_with_reference_runs compares with the mean of three runs, lets three times the reference's own
    # scatter about that mean widen the bar where it is the larger quantity, applies the same rule to the relative L2
    # error (which small entries cannot hide behind the largest one), and logs every figure.
    compare_with_reference_runs("config2_full", ours["grads"], [r["grads"] for r in refs], GRAD_KEYS)


@pytest.mark.skipif(REF is None, reason="baseline/_ref reference extension not built")
def test_config3_full_size_feature_only_training_against_reference():
    """BASELINE configs[2] at size: 500 k Gaussians, 1440x1080, language-feature-only training as train.py:96-99 runs
    it — only `language_feature` requires grad (scene/gaussian_model.py:212-217), the upstream gradient is the masked
    L1's (Bernoulli mask), no colour gradient."""
    scene, cams = syn.make_config(3)
    cam = cams[0]
    _, dfeat = syn.make_grads(cam.image_width, cam.image_height, 2003, feature_only=True)
    kw = dict(grads=(None, dfeat), requires=("language_feature",))
    ours = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, **kw)
    refs = [run_module(REF.GaussianRasterizer, REF.GaussianRasterizationSettings, scene, cam, **kw) for _ in range(3)]
    assert torch.equal(ours["radii"], refs[0]["radii"])
    assert (ours["color"] - refs[0]["color"]).abs().max().item() < 1e-5
    assert (ours["feature"] - refs[0]["feature"]).abs().max().item() < 1e-5
    for k in ("means3D", "shs", "opacities", "scales", "rotations"):
        assert ours["grads"][k] is None and refs[0]["grads"][k] is None
    compare_with_reference_runs("config3_full_feature_only", ours["grads"], [r["grads"] for r in refs],
                                ("language_feature", "means2D"))


def _multi_view_step(GR, GS, scene, cams, dev, factored=False):
    """One config-4 step: forward + backward of every view, gradients summed over the views (autograd accumulation,
    or this repo's factored SH path: per-view 12-byte factors, dL_dsh formed once)."""
    from helpers import settings_for
    names = ("means3D", "shs", "opacities", "scales", "rotations", "language_feature")
    params = {k: scene[k].to(dev).requires_grad_(True) for k in names}
    means2D = torch.zeros_like(params["means3D"], requires_grad=True)

    def run():
        for i, cam in enumerate(cams):
            dcol, dfeat = [t.to(dev) for t in syn.make_grads(cam.image_width, cam.image_height, 2400 + i)]
            rast = GR(raster_settings=settings_for(cam, dev, cls=GS))
            color, feat, radii = rast(means3D=params["means3D"], means2D=means2D, shs=params["shs"],
                                      language_feature_precomp=params["language_feature"],
                                      opacities=params["opacities"], scales=params["scales"],
                                      rotations=params["rotations"])
            torch.autograd.backward([color, feat], [dcol, dfeat])
            del color, feat, radii, dcol, dfeat

    if factored:
        from langsplat_b200 import dist as D
        with D.factored_sh_gradients() as factors:
            run()
        assert params["shs"].grad is None and len(factors) == len(cams)
        params["shs"].grad = D.finish_sh_gradients(factors, params["means3D"], params["shs"].shape[1])
    else:
        run()
    torch.cuda.synchronize()
    out = {k: p.grad.detach() for k, p in params.items()}
    out["means2D"] = means2D.grad.detach()
    return out


@pytest.mark.skipif(REF is None, reason="baseline/_ref reference extension not built")
def test_config4_full_size_eight_views_against_reference():
    """BASELINE configs[3] at size: 3 M Gaussians, 8 cameras per step, parameter gradients summed over the views — the
    dense path (autograd sums the per-view dL_dsh) and the factored SH path, each against the reference extension's
    autograd sum over the same eight views."""
    dev = torch.device("cuda:0")
    scene, cams = syn.make_config(4, n_cameras=8)
    refs = [_multi_view_step(REF.GaussianRasterizer, REF.GaussianRasterizationSettings, scene, cams, dev)
            for _ in range(2)]
    truth = {}

    def f64():  # evaluated at most once, and only if an element misses the bar (about a minute of host time)
        if not truth:
            truth.update(oracle_step_f64(scene, cams, 2400))
        return truth

    dense = _multi_view_step(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cams, dev)
    compare_with_reference_runs("config4_full_dense", dense, refs, GRAD_KEYS, arbiter=f64)
    del dense
    fact = _multi_view_step(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cams, dev, factored=True)
    compare_with_reference_runs("config4_full_factored_sh", fact, refs, GRAD_KEYS, arbiter=f64)


def test_config5_forward_structure_and_reference():
    """6 M Gaussians at 3840x2160 (the render-only sweep): the widest packed sort word (tile 15 + depth range + id 23
    bits)."""
    scene, cams = syn.make_config(5)
    cam = cams[0]
    color, feat, radii, state = cuda_forward_state(scene, cam)
    arr = R.debug_arrays(state)
    _check_structure(arr, state, cam.image_width, cam.image_height)
    if REF is None:
        pytest.skip("baseline/_ref not built: structure checked only")
    ref = ref_forward_buffers(REF, scene, cam)
    assert state["num_rendered"] == ref["num_rendered"]
    np.testing.assert_array_equal(_np(radii), _np(ref["radii"]))
    np.testing.assert_array_equal(_np(arr["point_list"]).view(np.uint32), ref["point_list"])
    np.testing.assert_array_equal(_np(arr["ranges"]).view(np.uint32), ref["ranges"])
    assert (color - ref["color"]).abs().max().item() < 1e-5
    assert (feat - ref["feature"]).abs().max().item() < 1e-5


def test_capacity_retry_gives_identical_results(monkeypatch):
    """A too-small instance list makes lsr_forward return LSR_ERR_CAPACITY with the needed count; the retry must
    reproduce the one-shot result exactly."""
    scene, cams = syn.make_config(1)
    cam = cams[0]
    base = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam)
    monkeypatch.setattr(R, "_last_R", {})
    monkeypatch.setattr(R, "_INITIAL_CAPACITY_FLOOR", 256)
    monkeypatch.setattr(R, "_CAPACITY_PER_GAUSSIAN", 0)
    monkeypatch.setattr(R, "_capacity_retries", 0)
    again = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam)
    assert R._capacity_retries >= 1
    assert torch.equal(base["color"], again["color"]) and torch.equal(base["radii"], again["radii"])
    assert torch.equal(base["feature"], again["feature"])
