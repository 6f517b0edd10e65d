"""Fused Adam step: the oracle's float64 restatement against golden vectors from torch.optim.Adam
configured like the reference (tests/golden/make_golden_adam.py), then — on the GPU — langsplat_b200.optim.FusedAdam
(lsr_adam_step) against the same vectors, against torch.optim.Adam on the device for more steps, the `visible`
(sparse) variant against the oracle, and optimizer-state compatibility with the reference's densification code.
Tolerance: an update is ~lr in size and every step rounds the float32 parameter once, so parameters are held to
1e-5 * lr + (steps + 1) ulp of the parameter; moments to 2e-6 relative."""
import os

import numpy as np
import pytest
import torch

from helpers import orc

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "adam.npz"))
NAMES = [str(n) for n in GOLD["names"]]
LRS = dict(zip(NAMES, (float(x) for x in GOLD["lrs"])))


def _close_param(got, want, lr, steps=1):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    ulp = np.spacing(np.maximum(np.abs(want), np.abs(got)).astype(np.float32)).astype(np.float64)
    tol = 1e-5 * lr + (steps + 1) * ulp
    assert np.all(np.abs(got - want) <= tol), float(np.abs(got - want).max())


def _close_rel(got, want, rel=2e-6):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    # exp_avg is a signed running sum: where its terms cancel the error scales with the terms, not the result
    floor = 0.1 * rel * float(np.abs(want).max(initial=0.0))
    assert np.all(np.abs(got - want) <= rel * np.abs(want) + floor + 1e-37), float(np.abs(got - want).max())


def test_oracle_matches_torch_adam_vectors():
    for k in NAMES:
        p, m, v = GOLD[f"{k}.p0"].astype(np.float64), 0.0 * GOLD[f"{k}.p0"], 0.0 * GOLD[f"{k}.p0"]
        for step in (1, 2, 3):
            p, m, v = orc.adam_step(p, GOLD[f"{k}.g{step}"], m, v, step, LRS[k])
            _close_param(p, GOLD[f"{k}.p{step}"], LRS[k], step)
            _close_rel(m, GOLD[f"{k}.m{step}"])
            _close_rel(v, GOLD[f"{k}.v{step}"])


@pytest.mark.gpu
def test_gpu_matches_torch_adam_vectors():
    from langsplat_b200.optim import FusedAdam
    dev = torch.device("cuda:0")
    params = {k: torch.nn.Parameter(torch.from_numpy(GOLD[f"{k}.p0"]).to(dev)) for k in NAMES}
    opt = FusedAdam([{"params": [params[k]], "lr": LRS[k], "name": k} for k in NAMES], lr=0.0, eps=1e-15)
    from langsplat_b200 import _lib
    before = _lib.load().lsr_launch_count()
    for step in (1, 2, 3):
        for k in NAMES:
            params[k].grad = torch.from_numpy(GOLD[f"{k}.g{step}"]).to(dev)
        opt.step()
        opt.zero_grad(set_to_none=True)
        for k in NAMES:
            st = opt.state[params[k]]
            _close_param(params[k].detach().cpu().numpy(), GOLD[f"{k}.p{step}"], LRS[k], step)
            _close_rel(st["exp_avg"].cpu().numpy(), GOLD[f"{k}.m{step}"])
            _close_rel(st["exp_avg_sq"].cpu().numpy(), GOLD[f"{k}.v{step}"])
            assert float(st["step"]) == step
    assert _lib.load().lsr_launch_count() - before == 3  # one launch per step for all six groups


@pytest.mark.gpu
def test_gpu_long_run_sparse_variant_and_state_compatibility():
    from langsplat_b200 import LsrError
    from langsplat_b200.optim import FusedAdam
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(9)
    P = 50021
    shapes = {"xyz": (P, 3), "f_rest": (P, 15, 3), "opacity": (P, 1), "rotation": (P, 4)}
    lrs = {"xyz": 1.6e-4, "f_rest": 1.25e-4, "opacity": 0.05, "rotation": 1e-3}
    init = {k: torch.randn(s, generator=g) for k, s in shapes.items()}
    ours = {k: torch.nn.Parameter(init[k].to(dev)) for k in shapes}
    theirs = {k: torch.nn.Parameter(init[k].to(dev)) for k in shapes}
    o1 = FusedAdam([{"params": [ours[k]], "lr": lrs[k], "name": k} for k in shapes], lr=0.0, eps=1e-15)
    o2 = torch.optim.Adam([{"params": [theirs[k]], "lr": lrs[k], "name": k} for k in shapes], lr=0.0, eps=1e-15)
    for step in range(1, 21):
        for k in shapes:
            grad = (torch.randn(shapes[k], generator=g) * 0.01).to(dev)
            ours[k].grad, theirs[k].grad = grad, grad.clone()
        o1.step()
        o2.step()
    for k in shapes:  # 20 steps of ~lr each: the two float32 implementations stay within a few ulp of an update
        _close_param(ours[k].detach().cpu().numpy(), theirs[k].detach().cpu().numpy(), lrs[k], 20)
        _close_rel(o1.state[ours[k]]["exp_avg_sq"].cpu().numpy(), o2.state[theirs[k]]["exp_avg_sq"].cpu().numpy(), 1e-5)
    # state_dict round trip into torch.optim.Adam and back (the reference restores checkpoints this way)
    o2.load_state_dict(o1.state_dict())
    o1.load_state_dict(o2.state_dict())
    # sparse variant against the oracle: only visible rows move, moments included
    vis = (torch.rand(P, generator=g) < 0.3)
    snap = {k: (ours[k].detach().cpu().numpy().copy(), o1.state[ours[k]]["exp_avg"].cpu().numpy().copy(),
                o1.state[ours[k]]["exp_avg_sq"].cpu().numpy().copy()) for k in shapes}
    grads = {k: (torch.randn(shapes[k], generator=g) * 0.01) for k in shapes}
    for k in shapes:
        ours[k].grad = grads[k].to(dev)
    o1.step(visible=vis.to(dev))
    for k in shapes:
        p, m, v = orc.adam_step(*snap[k][:1], grads[k].numpy(), snap[k][1], snap[k][2], 21, lrs[k],
                                visible=vis.numpy())
        _close_param(ours[k].detach().cpu().numpy(), p, lrs[k])
        _close_rel(o1.state[ours[k]]["exp_avg"].cpu().numpy(), m, 1e-5)
        hidden = ~vis.numpy()
        assert np.array_equal(ours[k].detach().cpu().numpy()[hidden], snap[k][0][hidden])
        assert np.array_equal(o1.state[ours[k]]["exp_avg_sq"].cpu().numpy()[hidden], snap[k][2][hidden])
    # the reference's densification edits optimizer state in place (scene/gaussian_model.py:300-331): prune rows
    keep = torch.rand(P, generator=g) < 0.5
    for group in o1.param_groups:
        old = group["params"][0]
        st = o1.state.pop(old)
        st["exp_avg"], st["exp_avg_sq"] = st["exp_avg"][keep.to(dev)], st["exp_avg_sq"][keep.to(dev)]
        new = torch.nn.Parameter(old.detach()[keep.to(dev)].contiguous())
        group["params"][0] = new
        o1.state[new] = st
        new.grad = torch.ones_like(new)
    o1.step()
    assert all(float(o1.state[g_["params"][0]]["step"]) == 22 for g_ in o1.param_groups)
    with pytest.raises(LsrError):
        bad = torch.nn.Parameter(torch.zeros(4, 3))
        bad.grad = torch.ones(4, 3)
        FusedAdam([bad], lr=0.1).step()  # CPU parameter: no CPU path
