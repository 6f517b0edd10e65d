"""The reference's own caller, UNMODIFIED, on top of this repo's rasterizer (north star: "render.py and train.py run
unchanged").

`gaussian_renderer/__init__.py` (render(), :19-115), `scene/gaussian_model.py` (GaussianModel: activations,
training_setup :191-227), `scene/cameras.py` (Camera: the transposed, NON-CONTIGUOUS world_view_transform of :54) and
`utils/loss_utils.py` are imported byte for byte from the reference checkout (/root/reference where it exists, else the
copy `baseline/stage_ref_caller.sh` stages under baseline/_ref/langsplat_caller — git-ignored like the built reference
extension, travels to the GPU box).  Only the two packages the reference needs for file I/O and initialisation and
that are not in this image (`plyfile`, `simple_knn._C`) are stubbed in sys.modules; neither is touched by the calls
below.

Each case runs one train.py:84-138-shaped iteration twice — render(), the loss of train.py:96-104, backward(),
optimizer.step() of the reference's own Adam groups — once with `import diff_gaussian_rasterization` resolving to this
repo's alias package (found through sys.path, nothing injected) and once resolving to the reference CUDA extension
(baseline/_ref), and compares images, radii, every parameter gradient, viewspace_points.grad and the updated
parameters.
"""
import importlib
import math
import os
import sys
import types

import numpy as np
import pytest
import torch

from helpers import ROOT, load_reference, rel_err, syn

pytestmark = pytest.mark.gpu

REF = load_reference()
_CALLER_DIRS = ("/root/reference", os.path.join(ROOT, "baseline", "_ref", "langsplat_caller"))
CALLER = next((d for d in _CALLER_DIRS if os.path.exists(os.path.join(d, "gaussian_renderer", "__init__.py"))), None)
_TOP = ("gaussian_renderer", "scene", "utils", "arguments", "diff_gaussian_rasterization", "plyfile", "simple_knn")


class _Caller:
    """The reference's caller modules imported with `diff_gaussian_rasterization` bound to one rasterizer."""

    def __init__(self, rasterizer_module):
        saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k.split(".")[0] in _TOP}
        ply = types.ModuleType("plyfile")
        ply.PlyData = ply.PlyElement = type("_Absent", (), {})
        knn, knn_c = types.ModuleType("simple_knn"), types.ModuleType("simple_knn._C")
        knn_c.distCUDA2 = lambda *_a, **_k: (_ for _ in ()).throw(RuntimeError("simple_knn is stubbed"))
        knn._C = knn_c
        sys.modules.update({"plyfile": ply, "simple_knn": knn, "simple_knn._C": knn_c})
        if rasterizer_module is not None:
            sys.modules["diff_gaussian_rasterization"] = rasterizer_module
        sys.path.insert(0, CALLER)
        try:
            self.renderer = importlib.import_module("gaussian_renderer")
            self.model = importlib.import_module("scene.gaussian_model")
            self.cameras = importlib.import_module("scene.cameras")
            self.loss = importlib.import_module("utils.loss_utils")
            self.general = importlib.import_module("utils.general_utils")
            self.rasterizer = sys.modules["diff_gaussian_rasterization"]
            for m in (self.renderer, self.model, self.cameras, self.loss):
                assert os.path.realpath(m.__file__).startswith(os.path.realpath(CALLER)), m.__file__
        finally:
            sys.path.remove(CALLER)
            for k in list(sys.modules):
                if k.split(".")[0] in _TOP:
                    del sys.modules[k]
            sys.modules.update(saved)


def _ours():
    c = _Caller(None)  # resolved by the import system: the alias package at the repo root
    assert os.path.realpath(c.rasterizer.__file__).startswith(os.path.join(os.path.realpath(ROOT),
                                                                           "diff_gaussian_rasterization"))
    import langsplat_b200
    assert c.rasterizer.GaussianRasterizer is langsplat_b200.GaussianRasterizer
    return c


def _training_args(include_feature):
    # arguments/__init__.py:74-94 (OptimizationParams defaults)
    return types.SimpleNamespace(include_feature=include_feature, percent_dense=0.01, position_lr_init=0.00016,
                                 position_lr_final=0.0000016, position_lr_delay_mult=0.01, position_lr_max_steps=30_000,
                                 feature_lr=0.0025, opacity_lr=0.05, scaling_lr=0.005, rotation_lr=0.001,
                                 language_feature_lr=0.0025, lambda_dssim=0.2)


def _build(caller, scene, cam, include_feature):
    """A GaussianModel and a Camera of the reference's own classes holding the synthetic scene."""
    nn = torch.nn
    pc = caller.model.GaussianModel(3)
    d = {k: v.cuda() for k, v in scene.items()}
    pc._xyz = nn.Parameter(d["means3D"].clone())
    pc._features_dc = nn.Parameter(d["shs"][:, 0:1, :].clone().contiguous())
    pc._features_rest = nn.Parameter(d["shs"][:, 1:, :].clone().contiguous())
    pc._scaling = nn.Parameter(torch.log(d["scales"]))
    pc._rotation = nn.Parameter(d["rotations"] * 1.7)  # un-normalised, as stored
    pc._opacity = nn.Parameter(caller.general.inverse_sigmoid(d["opacities"].clamp(1e-4, 1 - 1e-4)))
    pc.active_sh_degree = 3
    pc.spatial_lr_scale = 1.0
    if include_feature:
        pc._language_feature = nn.Parameter(d["language_feature"] * 0.8)
    pc.training_setup(_training_args(include_feature))
    V = cam.viewmatrix.double().t().numpy()  # world -> view
    Rm, T = V[:3, :3].T.copy(), V[:3, 3].copy()  # the colmap convention getWorld2View2 expects
    g = torch.Generator().manual_seed(4242)
    image = torch.rand(3, cam.image_height, cam.image_width, generator=g)
    camera = caller.cameras.Camera(colmap_id=0, R=Rm, T=T, FoVx=2 * math.atan(cam.tanfovx), FoVy=2 * math.atan(cam.tanfovy),
                                   image=image, gt_alpha_mask=None, image_name="synthetic", uid=0)
    assert not camera.world_view_transform.is_contiguous()  # scene/cameras.py:54
    gt_feat = torch.nn.functional.normalize(torch.rand(3, cam.image_height, cam.image_width, generator=g), dim=0).cuda()
    mask = (torch.rand(1, cam.image_height, cam.image_width, generator=g) < 0.8).cuda()  # bool, as seg != -1
    return pc, camera, gt_feat, mask


def _iteration(caller, scene, cam, include_feature, pipe_kw=None):
    pc, camera, gt_feat, mask = _build(caller, scene, cam, include_feature)
    opt = _training_args(include_feature)
    pipe = types.SimpleNamespace(debug=False, compute_cov3D_python=False, convert_SHs_python=False)
    for k, v in (pipe_kw or {}).items():
        setattr(pipe, k, v)
    bg = torch.tensor([0.0, 0.0, 0.0] if include_feature else [1.0, 1.0, 1.0], device="cuda")
    pkg = caller.renderer.render(camera, pc, pipe, bg, opt)
    image, feat = pkg["render"], pkg["language_feature_image"]
    if include_feature:  # train.py:96-99
        loss = caller.loss.l1_loss(feat * mask, gt_feat * mask)
    else:  # train.py:100-103
        gt_image = camera.original_image.cuda()
        ll1 = caller.loss.l1_loss(image, gt_image)
        loss = (1.0 - opt.lambda_dssim) * ll1 + opt.lambda_dssim * (1.0 - caller.loss.ssim(image, gt_image))
    loss.backward()
    names = ("_xyz", "_features_dc", "_features_rest", "_scaling", "_rotation", "_opacity", "_language_feature")
    grads = {n: getattr(pc, n).grad.detach().clone() for n in names
             if getattr(pc, n) is not None and getattr(pc, n).grad is not None}
    vsp = pkg["viewspace_points"].grad
    before = {n: getattr(pc, n).detach().clone() for n in grads}
    pc.optimizer.step()  # train.py:136-138
    pc.optimizer.zero_grad(set_to_none=True)
    delta = {n: (getattr(pc, n).detach() - before[n]) for n in grads}
    torch.cuda.synchronize()
    return dict(image=image.detach(), feature=feat.detach(), radii=pkg["radii"], visible=pkg["visibility_filter"],
                loss=float(loss.item()), grads=grads, viewspace=None if vsp is None else vsp.detach().clone(),
                delta=delta)


needs = pytest.mark.skipif(CALLER is None or REF is None,
                           reason="needs the staged reference caller (baseline/stage_ref_caller.sh) and baseline/_ref")


def _compare(a, b, include_feature):
    assert torch.equal(a["radii"], b["radii"]) and torch.equal(a["visible"], b["visible"])
    assert (a["image"] - b["image"]).abs().max().item() < 1e-5
    if include_feature:
        assert (a["feature"] - b["feature"]).abs().max().item() < 1e-5
    assert abs(a["loss"] - b["loss"]) <= 1e-6 * max(1.0, abs(b["loss"]))
    assert set(a["grads"]) == set(b["grads"]) and a["grads"], "the same parameters must receive gradients"
    for k in b["grads"]:
        err = rel_err(a["grads"][k], b["grads"][k])
        assert err < 1e-4, f"{k}.grad: rel err {err:.3e}"
    assert (a["viewspace"] is None) == (b["viewspace"] is None)
    if b["viewspace"] is not None:
        err = rel_err(a["viewspace"], b["viewspace"])
        assert err < 1e-4, f"viewspace_points.grad: rel err {err:.3e}"
    for k in b["delta"]:  # the reference's own Adam groups, first step: lr * sign(grad) wherever |grad| >> eps
        big = b["grads"][k].abs() > 1e-4 * b["grads"][k].abs().max()
        agree = (torch.sign(a["delta"][k][big]) == torch.sign(b["delta"][k][big])).float().mean().item()
        assert agree > 0.999, f"{k}: parameter updates agree in sign on {agree:.5f} of the significant entries"


@needs
@pytest.mark.parametrize("include_feature", [True, False])
def test_reference_render_and_train_iteration_run_unchanged(include_feature):
    scene, cams = syn.make_config(1)
    ours = _iteration(_ours(), scene, cams[0], include_feature)
    ref = _iteration(_Caller(REF), scene, cams[0], include_feature)
    if include_feature:  # gaussian_model.py:212-217: only the language feature trains
        assert set(ours["grads"]) == {"_language_feature"}
    else:
        assert set(ours["grads"]) == {"_xyz", "_features_dc", "_features_rest", "_scaling", "_rotation", "_opacity"}
    _compare(ours, ref, include_feature)


@needs
@pytest.mark.parametrize("pipe_kw", [dict(convert_SHs_python=True), dict(compute_cov3D_python=True)])
def test_reference_python_side_branches(pipe_kw):
    """render() with colours / 3D covariances precomputed in Python (gaussian_renderer/__init__.py:62-83): the
    colors_precomp and cov3D_precomp inputs of the rasterizer, gradients flowing back through the reference's own
    eval_sh / build_scaling_rotation."""
    scene, cams = syn.make_config(1, P=6000)
    ours = _iteration(_ours(), scene, cams[0], False, pipe_kw)
    ref = _iteration(_Caller(REF), scene, cams[0], False, pipe_kw)
    _compare(ours, ref, False)


@needs
def test_reference_render_at_half_hd():
    """A larger frame (200 k Gaussians, 960x540) through the same unmodified caller, feature mode."""
    scene, cams = syn.make_config(2, P=200_000, W=960, H=540, n_cameras=2)
    ours = _iteration(_ours(), scene, cams[1], True)
    ref = _iteration(_Caller(REF), scene, cams[1], True)
    _compare(ours, ref, True)


@pytest.mark.skipif(CALLER is None, reason="reference caller not staged")
def test_staged_caller_is_byte_identical():
    sums = os.path.join(CALLER, "SHA256SUMS")
    if not os.path.exists(sums):
        pytest.skip("running against /root/reference itself")
    import hashlib
    for line in open(sums):
        digest, name = line.split()
        assert hashlib.sha256(open(os.path.join(CALLER, name), "rb").read()).hexdigest() == digest, name
