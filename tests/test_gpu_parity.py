"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.

Bars (BASELINE.json north_star): bit-exact radii / tile keys / sorted point list / tile ranges; images within 1e-5
max-abs; gradients within 1e-4 relative (max-abs error over max-abs value, because the reference itself sums with
unordered atomics).
"""
import numpy as np
import pytest
import torch

from helpers import (R, cuda_forward_state, oracle_forward, orc, rel_err, run_module, settings_for, syn)

pytestmark = pytest.mark.gpu

IMG_TOL = 1e-5
GRAD_TOL = 1e-4


def _np(t):
    return t.detach().cpu().numpy()


def _compare_forward(scene, cam, **kw):
    fw = oracle_forward(scene, cam, **kw)
    color, feat, radii, state = cuda_forward_state(scene, cam, **kw)
    arr = R.debug_arrays(state)
    Rn = fw["num_rendered"]
    assert state["num_rendered"] == Rn
    vis = fw["radii"] > 0
    # ---- bit-exact integer / key quantities ----
    np.testing.assert_array_equal(_np(radii), fw["radii"])
    np.testing.assert_array_equal(_np(arr["tiles_touched"]).view(np.uint32), fw["tiles_touched"])
    # the CUDA path lays the instances out in depth order (binning.cu): its prefix sum is the reference's
    # point_offsets (rasterizer_impl.cu:280) taken over the stable depth ordering of the Gaussians
    if Rn > 0:
        order = _np(arr["depth_order"]).view(np.uint32)
        assert np.array_equal(np.sort(order), np.arange(order.size, dtype=np.uint32))
        touched = fw["tiles_touched"] > 0
        dbits = fw["depths"].view(np.uint32).astype(np.int64)
        dkey = np.where(touched, dbits - dbits[touched].min(), 0)  # Gaussians without tiles carry depth field 0
        np.testing.assert_array_equal(order, np.argsort(dkey, kind="stable").astype(np.uint32))
        np.testing.assert_array_equal(_np(arr["point_offsets"]).view(np.uint32),
                                      np.cumsum(fw["tiles_touched"][order], dtype=np.uint32))
    rec = _np(arr["rec"])
    np.testing.assert_array_equal(rec[vis, 0:2].view(np.uint32), fw["means2D"][vis].view(np.uint32))
    np.testing.assert_array_equal(_np(arr["depths"])[vis].view(np.uint32), fw["depths"][vis].view(np.uint32))
    np.testing.assert_array_equal(_np(arr["keys"]).view(np.uint64), fw["keys"])
    np.testing.assert_array_equal(_np(arr["point_list"]).view(np.uint32), fw["point_list"])
    np.testing.assert_array_equal(_np(arr["ranges"]).view(np.uint32), fw["ranges"])
    # conic: same rounding chain -> bit-exact too
    np.testing.assert_array_equal(rec[vis][:, [2, 3, 4]].view(np.uint32),
                                  fw["conic_opacity"][vis][:, 0:3].view(np.uint32))
    if kw.get("cov3D_precomp") is None:
        np.testing.assert_array_equal(_np(arr["cov3D"])[vis].view(np.uint32), fw["cov3D"][vis].view(np.uint32))
    if kw.get("colors_precomp") is None:
        np.testing.assert_allclose(rec[vis][:, 6:9], fw["rgb"][vis], atol=2e-6, rtol=0)
    # ---- images ----
    np.testing.assert_array_equal(_np(arr["n_contrib"]).view(np.uint32), fw["n_contrib"])
    np.testing.assert_allclose(_np(arr["final_T"]), fw["final_T"], atol=IMG_TOL, rtol=0)
    np.testing.assert_allclose(_np(color), fw["color"], atol=IMG_TOL, rtol=0)
    if kw.get("include_feature", True):
        np.testing.assert_allclose(_np(feat), fw["feature"], atol=IMG_TOL, rtol=0)
    return fw, state


def test_forward_config1_bit_exact_and_images():
    scene, cams = syn.make_config(1)
    fw, _ = _compare_forward(scene, cams[0])
    assert fw["num_rendered"] > 50000


@pytest.mark.parametrize("W,H", [(250, 130), (17, 16), (640, 360)])
def test_forward_ragged_sizes(W, H):
    scene, cams = syn.make_config(1, P=5000, W=W, H=H)
    _compare_forward(scene, cams[0])


@pytest.mark.parametrize("deg", [0, 1, 2])
def test_forward_lower_sh_degrees(deg):
    scene, cams = syn.make_config(1, P=4096, W=128, H=128)
    _compare_forward(scene, cams[0], sh_degree=deg)


def test_forward_background_and_no_feature():
    scene, cams = syn.make_config(1, P=4096, W=128, H=96)
    _compare_forward(scene, cams[0], bg=(0.2, 0.5, 1.0), include_feature=False)


def test_forward_scale_modifier():
    scene, cams = syn.make_config(1, P=4096, W=128, H=96)
    _compare_forward(scene, cams[0], scale_modifier=1.7)


def test_forward_all_culled_and_empty():
    scene, cams = syn.make_config(1, P=1000, W=64, H=64)
    scene = dict(scene)
    scene["means3D"] = scene["means3D"] + torch.tensor([0.0, 0.0, -50.0])  # all behind the camera
    fw, state = _compare_forward(scene, cams[0])
    assert fw["num_rendered"] == 0 and state["num_rendered"] == 0
    # P == 0: the reference returns zero images (rasterize_points.cu:70-91)
    empty = {k: v[:0] for k, v in scene.items()}
    out = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, empty, cams[0], bg=(1.0, 1.0, 1.0))
    assert out["color"].abs().max().item() == 0 and out["radii"].numel() == 0


def _oracle_grads(scene, cam, dcol, dfeat, **kw):
    fw = oracle_forward(scene, cam, **kw)
    H, W = cam.image_height, cam.image_width
    g = orc.backward(fw, dcol if dcol is not None else np.zeros((3, H, W), np.float32), dfeat)
    return fw, g


def _check_grads(cu, g, keys, tol=GRAD_TOL):
    for k_cu, k_or in keys:
        a, b = cu[k_cu], g[k_or]
        assert a is not None, k_cu
        b = torch.from_numpy(b).reshape(a.shape)
        err = rel_err(a, b)
        assert err < tol, f"{k_cu}: rel err {err:.3e}"


ALL_KEYS = [("means3D", "means3D"), ("means2D", "means2D"), ("shs", "shs"), ("language_feature", "language_feature"),
            ("opacities", "opacities"), ("scales", "scales"), ("rotations", "rotations")]


def test_backward_config1_all_gradients():
    scene, cams = syn.make_config(1)
    cam = cams[0]
    dcol, dfeat = syn.make_grads(cam.image_width, cam.image_height, 2001)
    _, g = _oracle_grads(scene, cam, dcol.numpy(), dfeat.numpy())
    out = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, grads=(dcol, dfeat))
    _check_grads(out["grads"], g, ALL_KEYS)


def test_backward_with_background_ragged():
    scene, cams = syn.make_config(1, P=6000, W=200, H=120)
    cam = cams[0]
    dcol, dfeat = syn.make_grads(cam.image_width, cam.image_height, 7)
    _, g = _oracle_grads(scene, cam, dcol.numpy(), dfeat.numpy(), bg=(0.3, 0.6, 0.9))
    out = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, grads=(dcol, dfeat),
                     bg=(0.3, 0.6, 0.9))
    _check_grads(out["grads"], g, ALL_KEYS)


def test_backward_feature_only_mode():
    """LangSplat feature training (train.py:96-99, gaussian_model.py:212-217): only language_feature (and means2D)
    require grad and the loss never touches the colour image."""
    scene, cams = syn.make_config(3, P=20000, W=360, H=270)
    cam = cams[0]
    _, dfeat = syn.make_grads(cam.image_width, cam.image_height, 2003, feature_only=True)
    _, g = _oracle_grads(scene, cam, None, dfeat.numpy())
    out = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, grads=(None, dfeat),
                     requires=("language_feature",))
    _check_grads(out["grads"], g, [("language_feature", "language_feature"), ("means2D", "means2D")])
    for k in ("means3D", "shs", "opacities", "scales", "rotations"):
        assert out["grads"][k] is None


def test_backward_no_feature_rgb_mode():
    scene, cams = syn.make_config(1, P=8000, W=160, H=160)
    cam = cams[0]
    dcol, _ = syn.make_grads(cam.image_width, cam.image_height, 11)
    _, g = _oracle_grads(scene, cam, dcol.numpy(), None, include_feature=False)
    out = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, grads=(dcol, None),
                     include_feature=False, requires=("means3D", "shs", "opacities", "scales", "rotations"))
    _check_grads(out["grads"], g, [k for k in ALL_KEYS if k[0] != "language_feature"])


def test_precomputed_colour_and_covariance_branches():
    """The reference's own cross-checks (--convert_SHs_python / --compute_cov3D_python, gaussian_renderer:56-80):
    feeding the oracle's rgb / cov3D as precomputed inputs must reproduce the SH / scale-rotation render."""
    scene, cams = syn.make_config(1, P=6000, W=160, H=120)
    cam = cams[0]
    fw = oracle_forward(scene, cam)
    base = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam)
    cp = torch.from_numpy(fw["rgb"].copy())
    c3 = torch.from_numpy(fw["cov3D"].copy())
    # rows of culled Gaussians are never written by the oracle; any finite value works there
    dcol, dfeat = syn.make_grads(cam.image_width, cam.image_height, 5)
    alt = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, use_colors_precomp=cp,
                     use_cov3D_precomp=c3, grads=(dcol, dfeat),
                     requires=("means3D", "colors_precomp", "cov3D_precomp", "opacities", "language_feature"))
    vis = torch.from_numpy(fw["radii"] > 0)
    assert torch.equal(alt["radii"].cpu()[vis], base["radii"].cpu()[vis])
    assert (alt["color"] - base["color"]).abs().max().item() < IMG_TOL
    assert (alt["feature"] - base["feature"]).abs().max().item() < IMG_TOL
    fw2 = oracle_forward(scene, cam, colors_precomp=fw["rgb"], cov3D_precomp=fw["cov3D"])
    g = orc.backward(fw2, dcol.numpy(), dfeat.numpy())
    _check_grads(alt["grads"], g, [("colors_precomp", "colors"), ("cov3D_precomp", "cov3D"), ("means3D", "means3D"),
                                   ("opacities", "opacities"), ("language_feature", "language_feature")])
    assert alt["grads"]["scales"] is None and alt["grads"]["shs"] is None


def test_mark_visible():
    scene, cams = syn.make_config(1, P=10000)
    cam = cams[0]
    scene["means3D"][::3] += torch.tensor([0.0, 0.0, -3.0])
    rast = R.GaussianRasterizer(R.GaussianRasterizationSettings(
        cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, torch.zeros(3, device="cuda"), 1.0,
        cam.viewmatrix.cuda(), cam.projmatrix.cuda(), 3, cam.campos.cuda(), False, False, True))
    got = rast.markVisible(scene["means3D"].cuda()).cpu().numpy()
    want = orc.mark_visible(scene["means3D"], cam.viewmatrix)
    np.testing.assert_array_equal(got, want)
    assert 0 < want.sum() < want.size


def test_api_errors_match_reference_messages():
    scene, cams = syn.make_config(1, P=64, W=32, H=32)
    cam = cams[0]
    d = {k: v.cuda() for k, v in scene.items()}
    from helpers import settings_for
    rast = R.GaussianRasterizer(settings_for(cam, "cuda"))
    m2 = torch.zeros_like(d["means3D"])
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        rast(d["means3D"], m2, d["opacities"], scales=d["scales"], rotations=d["rotations"])
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        rast(d["means3D"], m2, d["opacities"], shs=d["shs"])
    with pytest.raises(RuntimeError, match=r"means3D must have dimensions \(num_points, 3\)"):
        rast(d["means3D"][:, :2], m2, d["opacities"], shs=d["shs"], scales=d["scales"], rotations=d["rotations"],
             language_feature_precomp=d["language_feature"])


def test_non_contiguous_camera_matrices():
    """scene/cameras.py:54 hands the rasterizer a transposed (non-contiguous) view matrix."""
    scene, cams = syn.make_config(1, P=3000, W=96, H=96)
    cam = cams[0]
    base = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam)
    cam2 = syn.Camera(cam.image_width, cam.image_height, cam.tanfovx, cam.tanfovy,
                      cam.viewmatrix.t().contiguous().t(), cam.projmatrix.t().contiguous().t(), cam.campos)
    assert not cam2.viewmatrix.is_contiguous()
    alt = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam2)
    assert torch.equal(base["color"], alt["color"]) and torch.equal(base["radii"], alt["radii"])


def test_misaligned_input_views_are_repacked():
    """Contiguous views that start at an odd element of their storage (x[1:]): the kernels' 128-bit accesses need
    16-byte alignment, so the host layer re-packs such tensors instead of faulting; results are unchanged."""
    scene, cams = syn.make_config(1, P=3001, W=96, H=96)
    cam = cams[0]
    dcol, dfeat = syn.make_grads(cam.image_width, cam.image_height, 31)
    cut = {k: v[1:].contiguous() for k, v in scene.items()}
    base = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, cut, cam, grads=(dcol, dfeat))
    rs = settings_for(cam, "cuda")
    dev = {k: v.cuda() for k, v in scene.items()}
    t = {k: v[1:].requires_grad_(True) for k, v in dev.items()}  # views: storage offset of one row
    assert t["rotations"].data_ptr() % 16 == 0 and t["means3D"].data_ptr() % 16 != 0
    rot_odd = torch.cat([torch.zeros(1, device="cuda"), dev["rotations"][1:].reshape(-1)])[1:].view(-1, 4)
    assert rot_odd.data_ptr() % 16 == 4 and rot_odd.is_contiguous()
    rot_odd.requires_grad_(True)
    means2D = torch.zeros_like(t["means3D"], requires_grad=True)
    color, feat, radii = R.GaussianRasterizer(rs)(
        means3D=t["means3D"], means2D=means2D, shs=t["shs"], language_feature_precomp=t["language_feature"],
        opacities=t["opacities"], scales=t["scales"], rotations=rot_odd)
    ((color * dcol.cuda()).sum() + (feat * dfeat.cuda()).sum()).backward()
    torch.cuda.synchronize()
    assert torch.equal(color, base["color"]) and torch.equal(radii, base["radii"])
    assert rel_err(rot_odd.grad, base["grads"]["rotations"]) < 1e-5


def test_in_place_update_between_forward_and_backward_is_detected():
    """The reference saves its inputs with ctx.save_for_backward (RAST/diff_gaussian_rasterization/__init__.py:102), so
    autograd raises when one of them is modified in place before backward; same here."""
    scene, cams = syn.make_config(1, P=2000, W=64, H=64)
    cam = cams[0]
    t = {k: v.cuda().requires_grad_(True) for k, v in scene.items()}
    means2D = torch.zeros_like(t["means3D"], requires_grad=True)
    color, feat, radii = R.GaussianRasterizer(settings_for(cam, "cuda"))(
        means3D=t["means3D"], means2D=means2D, shs=t["shs"], language_feature_precomp=t["language_feature"],
        opacities=t["opacities"], scales=t["scales"], rotations=t["rotations"])
    with torch.no_grad():
        t["scales"].mul_(1.01)  # e.g. an optimizer step between the forwards and backwards of a multi-view step
    with pytest.raises(RuntimeError, match="modified by an inplace operation"):
        (color.sum() + feat.sum()).backward()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_second_device_in_the_same_process():
    """Function attributes (dynamic shared memory > 48 KB) are per device: forward + backward on cuda:1 after cuda:0."""
    scene, cams = syn.make_config(1, P=5000, W=128, H=128)
    cam = cams[0]
    dcol, dfeat = syn.make_grads(cam.image_width, cam.image_height, 32)
    a = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, device="cuda:0", grads=(dcol, dfeat))
    b = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, device="cuda:1", grads=(dcol, dfeat))
    assert torch.equal(a["color"].cpu(), b["color"].cpu()) and torch.equal(a["radii"].cpu(), b["radii"].cpu())
    for k in ("means3D", "shs", "scales", "rotations", "opacities", "language_feature"):
        assert rel_err(a["grads"][k].cpu(), b["grads"][k].cpu()) < 1e-5
