"""GPU: the forward without host synchronisation (lsr_forward_nosync, langsplat_b200.set_async_forward).

Same kernels as the synchronous path, but the instance count and the depth range stay on the device (the reference
reads the count back in the middle of every frame, rasterizer_impl.cu:284).  Checked here: results identical to the
synchronous forward bit for bit; an instance list that turns out too small NaN-fills the images and raises at the next
call instead of returning an incomplete render; forward + backward can be captured in a CUDA graph (which also proves
that the library makes no synchronising call) and replayed with new camera matrices."""
import numpy as np
import pytest
import torch

import langsplat_b200 as lsb
from helpers import R, rel_err, run_module, settings_for, syn

pytestmark = pytest.mark.gpu


def _np(t):
    return t.detach().cpu().numpy()


@pytest.fixture(autouse=True)
def _sync_default():
    yield
    lsb.set_async_forward(False, 1.5)
    for pend in R._pending.values():
        pend.clear()


def _forward_state(scene, cam, **kw):
    rs = settings_for(cam, "cuda", **kw)
    d = {k: v.cuda() for k, v in scene.items()}
    out = R._forward_impl(rs, d["means3D"], d["shs"], None, d["language_feature"], d["opacities"], d["scales"],
                          d["rotations"], None, keep_scratch=True)
    return out


@pytest.mark.parametrize("cfg,P,W,H", [(1, None, None, None), (2, 120_000, 960, 540), (1, 5, 64, 48)])
def test_async_forward_is_bit_identical(cfg, P, W, H):
    scene, cams = syn.make_config(cfg, P=P, W=W, H=H)
    cam = cams[0]
    c0, f0, r0, s0 = _forward_state(scene, cam)  # synchronous: also teaches the device the instance count
    a0 = R.debug_arrays(s0)
    assert s0["num_rendered"] is not None
    with lsb.async_forward():
        c1, f1, r1, s1 = _forward_state(scene, cam)
        assert s1["num_rendered"] is None and s1["async"].status is not None  # nothing was read back
        a1 = R.debug_arrays(s1)  # (waits for the status words)
    assert s1["num_rendered"] == s0["num_rendered"]
    assert torch.equal(c0, c1) and torch.equal(f0, f1) and torch.equal(r0, r1)
    for k in ("ranges", "n_contrib", "final_T", "point_list", "point_offsets", "depth_order", "keys"):
        assert torch.equal(a0[k], a1[k]), k


def test_async_gradients_match_the_synchronous_path():
    scene, cams = syn.make_config(1)
    cam = cams[0]
    dcol, dfeat = syn.make_grads(cam.image_width, cam.image_height, 77)
    base = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, grads=(dcol, dfeat))
    again = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, grads=(dcol, dfeat))
    with lsb.async_forward():
        out = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, grads=(dcol, dfeat))
        lsb.check_async_status()
    assert torch.equal(out["color"], base["color"]) and torch.equal(out["feature"], base["feature"])
    for k in ("means3D", "means2D", "shs", "language_feature", "opacities", "scales", "rotations"):
        noise = rel_err(again["grads"][k], base["grads"][k])  # unordered float atomics in the blend backward
        assert rel_err(out["grads"][k], base["grads"][k]) <= max(1e-6, 4.0 * noise), k


def test_overflow_poisons_the_images_and_raises_at_the_next_call():
    scene, cams = syn.make_config(1)
    cam = cams[0]
    base = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam)
    true_R = R._last_R[torch.cuda.current_device()]
    R._recent_R[torch.cuda.current_device()].clear()
    lsb.reserve_instances("cuda", 1000)  # far below the frame's instance count (capacity floor 64 Ki)
    assert true_R > (1 << 16)
    with lsb.async_forward(True, 1.0):
        out = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam)
        assert torch.isnan(out["color"]).all() and torch.isnan(out["feature"]).all()
        assert torch.equal(out["radii"], base["radii"])  # the per-Gaussian part of the frame is complete
        with pytest.raises(lsb.LsrError, match="instances but its list was sized for"):
            lsb.check_async_status()
        # the count is known now: the next frame gets a list that fits
        again = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam)
        lsb.check_async_status()
    assert torch.equal(again["color"], base["color"]) and torch.equal(again["feature"], base["feature"])


def test_overflow_is_reported_by_the_next_forward_without_an_explicit_check():
    scene, cams = syn.make_config(1)
    cam = cams[0]
    run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam)
    R._recent_R[torch.cuda.current_device()].clear()
    lsb.reserve_instances("cuda", 1000)
    with lsb.async_forward(True, 1.0):
        run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam)
        torch.cuda.synchronize()  # the status words have arrived in pinned memory
        with pytest.raises(lsb.LsrError, match="instances but its list was sized for"):
            run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam)


def test_forward_and_backward_replay_from_a_cuda_graph():
    """One training-step-shaped graph: forward, a loss, backward — replayed with another camera's matrices copied
    into the captured buffers.  Capturing fails if anything on the path synchronises."""
    scene, cams = syn.make_config(1, n_cameras=2)
    dev = torch.device("cuda")
    dcol, dfeat = [t.to(dev) for t in syn.make_grads(cams[0].image_width, cams[0].image_height, 5)]

    def eager(cam):
        return run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cam, grads=(dcol, dfeat))

    want = [eager(c) for c in cams]  # synchronous; also the capacity hint (largest of the two views)
    names = ("means3D", "shs", "language_feature", "opacities", "scales", "rotations")
    params = {k: scene[k].to(dev).requires_grad_(True) for k in names}
    means2D = torch.zeros_like(params["means3D"], requires_grad=True)
    view, proj, campos = (cams[0].viewmatrix.to(dev).clone(), cams[0].projmatrix.to(dev).clone(),
                          cams[0].campos.to(dev).clone())
    rs = settings_for(cams[0], dev)._replace(viewmatrix=view, projmatrix=proj, campos=campos)
    rast = R.GaussianRasterizer(rs)
    grads_out = {}

    def step():
        color, feat, radii = rast(means3D=params["means3D"], means2D=means2D, shs=params["shs"],
                                  language_feature_precomp=params["language_feature"], opacities=params["opacities"],
                                  scales=params["scales"], rotations=params["rotations"])
        loss = (color * dcol).sum() + (feat * dfeat).sum()
        gs = torch.autograd.grad(loss, [params[k] for k in names] + [means2D])
        return color, feat, radii, gs

    with lsb.async_forward():
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):  # warm-up outside the capture (allocations, lazy initialisation)
                step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            color, feat, radii, gs = step()
        for i, cam in enumerate(cams):
            view.copy_(cam.viewmatrix.to(dev)); proj.copy_(cam.projmatrix.to(dev)); campos.copy_(cam.campos.to(dev))
            graph.replay()
            torch.cuda.synchronize()
            assert torch.equal(color, want[i]["color"]) and torch.equal(feat, want[i]["feature"])
            assert torch.equal(radii, want[i]["radii"])
            got = dict(zip(names + ("means2D",), gs))
            for k in names + ("means2D",):
                assert rel_err(got[k], want[i]["grads"][k]) < 1e-4, (i, k)  # the path's gradient bar (unordered atomics)
        for pend in R._pending.values():  # (the capture's status stamp cannot be matched to a replay)
            pend.clear()


def test_graphed_step_helper():
    """langsplat_b200.graphs.GraphedStep: capture forward + fused L1 loss + backward once, replay per view."""
    from langsplat_b200.graphs import GraphedStep
    scene, cams = syn.make_config(1, n_cameras=2)
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(9)
    gts = [(torch.rand(3, c.image_height, c.image_width, generator=g).to(dev),
            torch.rand(3, c.image_height, c.image_width, generator=g).to(dev)) for c in cams]
    names = ("means3D", "shs", "language_feature", "opacities", "scales", "rotations")
    params = {k: scene[k].to(dev).requires_grad_(True) for k in names}
    means2D = torch.zeros_like(params["means3D"], requires_grad=True)

    def iteration(rast, gt_c, gt_f):
        color, feat, radii = rast(means3D=params["means3D"], means2D=means2D, shs=params["shs"],
                                  language_feature_precomp=params["language_feature"], opacities=params["opacities"],
                                  scales=params["scales"], rotations=params["rotations"])
        loss = lsb.l1_loss(color, gt_c) + lsb.l1_loss(feat, gt_f)
        return loss, torch.autograd.grad(loss, [params[k] for k in names])

    want = []
    for cam, (gt_c, gt_f) in zip(cams, gts):  # eager, synchronous
        loss, grads = iteration(R.GaussianRasterizer(settings_for(cam, dev)), gt_c, gt_f)
        want.append((loss.item(), [x.clone() for x in grads]))
    del loss, grads
    # fresh leaves for the captured iteration: autograd nodes created by the eager runs above belong to the default
    # stream, and a capture must not make the legacy stream depend on the capturing one
    params = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
    means2D = torch.zeros_like(params["means3D"], requires_grad=True)
    torch.cuda.synchronize()
    view, proj, campos = (cams[0].viewmatrix.to(dev).clone(), cams[0].projmatrix.to(dev).clone(),
                          cams[0].campos.to(dev).clone())
    gt_c, gt_f = gts[0][0].clone(), gts[0][1].clone()
    rast = R.GaussianRasterizer(settings_for(cams[0], dev)._replace(viewmatrix=view, projmatrix=proj, campos=campos))
    step = GraphedStep(lambda: iteration(rast, gt_c, gt_f))
    assert not lsb.rasterizer._AsyncConfig.enabled  # the switch is restored after the capture
    for i in (1, 0, 1):
        view.copy_(cams[i].viewmatrix.to(dev)); proj.copy_(cams[i].projmatrix.to(dev)); campos.copy_(cams[i].campos.to(dev))
        gt_c.copy_(gts[i][0]); gt_f.copy_(gts[i][1])
        loss, grads = step.replay()
        step.check()
        assert abs(loss.item() - want[i][0]) <= 1e-6 * abs(want[i][0])
        for got, ref in zip(grads, want[i][1]):
            assert rel_err(got, ref) < 1e-4
