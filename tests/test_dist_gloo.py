"""CPU, world_size 2 over gloo: the view-sharded step (BASELINE.json config 4).  Each rank rasterises
its share of the step's camera views against the replicated Gaussians (here with the CPU oracle standing in for the
kernels — the host-side sharding / reduction logic is what is under test), sums its views' gradients, and one
sum-allreduce leaves every rank with the gradient of the whole step."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import oracle_forward, orc, syn
from langsplat_b200 import dist as D

N_VIEWS = 4
NAMES = ("means3D", "shs", "opacities", "scales", "rotations", "language_feature")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _view_grads(scene, cam, seed):
    dcol, dfeat = syn.make_grads(cam.image_width, cam.image_height, seed)
    fw = oracle_forward(scene, cam)
    g = orc.backward(fw, dcol.numpy(), dfeat.numpy())
    stats = dict(norm=np.linalg.norm(g["means2D"][:, :2], axis=1), vis=(fw["radii"] > 0), radii=fw["radii"])
    return {k: torch.from_numpy(np.asarray(g[k], np.float32).reshape(scene[k].shape)) for k in NAMES}, stats


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        orc.set_num_threads(2)
        scene, cams = syn.make_config(4, P=1500, W=96, H=64, n_cameras=N_VIEWS)
        mine = D.views_for_rank(N_VIEWS, rank, world)
        acc = {k: torch.zeros_like(scene[k]) for k in NAMES}
        s_sum, s_cnt = torch.zeros(1500), torch.zeros(1500)
        s_max = torch.zeros(1500, dtype=torch.int32)
        for v in mine:
            g, st = _view_grads(scene, cams[v], 100 + v)
            for k in NAMES:
                acc[k] += g[k]
            s_sum += torch.from_numpy(st["norm"]).float() * torch.from_numpy(st["vis"])
            s_cnt += torch.from_numpy(st["vis"]).float()
            s_max = torch.maximum(s_max, torch.from_numpy(st["radii"]))
        handles = D.allreduce_gradients([acc[k] for k in NAMES] + [None], async_op=True)
        for h in handles:
            h.wait()
        D.reduce_view_statistics(s_sum, s_cnt, s_max)
        torch.save(dict(acc=acc, views=mine, s_sum=s_sum, s_cnt=s_cnt, s_max=s_max),
                   os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_views_for_rank_partitions_the_step():
    for world in (1, 2, 3, 4, 8):
        seen = sorted(v for r in range(world) for v in D.views_for_rank(8, r, world))
        assert seen == list(range(8))
    assert D.views_for_rank(8, 1, 4) == [1, 5]
    assert D.allreduce_gradients([torch.ones(3)]) == []  # no process group: a no-op, not an error


@pytest.mark.timeout(300)
def test_view_sharded_step_equals_serial_sum(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    r0 = torch.load(tmp_path / "rank0.pt")
    r1 = torch.load(tmp_path / "rank1.pt")
    assert sorted(r0["views"] + r1["views"]) == list(range(N_VIEWS))
    scene, cams = syn.make_config(4, P=1500, W=96, H=64, n_cameras=N_VIEWS)
    want = {k: torch.zeros_like(scene[k]) for k in NAMES}
    cnt = torch.zeros(1500)
    for v in range(N_VIEWS):
        g, st = _view_grads(scene, cams[v], 100 + v)
        for k in NAMES:
            want[k] += g[k]
        cnt += torch.from_numpy(st["vis"]).float()
    for k in NAMES:
        assert torch.equal(r0["acc"][k], r1["acc"][k]), k  # every rank ends with the same reduced gradient
        scale = want[k].abs().max().item()
        assert (r0["acc"][k] - want[k]).abs().max().item() <= 1e-5 * scale, k
    assert torch.equal(r0["s_cnt"], cnt) and torch.equal(r1["s_cnt"], cnt)
    assert torch.equal(r0["s_max"], r1["s_max"]) and r0["s_max"].max() > 0


def test_coalesce_merges_adjacent_gradient_slices():
    flat = torch.arange(100, dtype=torch.float32)
    a, b, c = flat[0:12].view(4, 3), flat[12:60].view(4, 4, 3), flat[60:64].view(4, 1)
    gap = flat[70:76].view(2, 3)
    other = torch.ones(5)
    out = D.coalesce([b, None, a, gap, c, other])
    sizes = sorted(t.numel() for t in out)
    assert sizes == [5, 6, 64]
    big = [t for t in out if t.numel() == 64][0]
    assert big.data_ptr() == flat.data_ptr() and torch.equal(big, flat[:64])
    big += 1.0  # the merged view aliases the original slices
    assert a[0, 0].item() == 1.0 and c[3, 0].item() == 64.0


# ---- factored SH gradients: the exchange (gather_sh_factors) and the rank-1 claim it rests on ----------------------
# SH real-basis constants as published with 3DGS (RAST cuda_rasterizer/auxiliary.h:23-40)
_C0, _C1 = 0.28209479177387814, 0.4886025119029199
_C2 = (1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396)
_C3 = (-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
       1.445305721320277, -0.5900435899266435)


def _sh_basis(d):
    """[P,16] degree-3 basis at unit directions d [P,3] (forward.cu:20-70 evaluates colour = sum_k basis_k * sh_k)."""
    x, y, z = d[:, 0], d[:, 1], d[:, 2]
    xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
    return np.stack([np.full_like(x, _C0), -_C1 * y, _C1 * z, -_C1 * x,
                     _C2[0] * xy, _C2[1] * yz, _C2[2] * (2 * zz - xx - yy), _C2[3] * xz, _C2[4] * (xx - yy),
                     _C3[0] * y * (3 * xx - yy), _C3[1] * xy * z, _C3[2] * y * (4 * zz - xx - yy),
                     _C3[3] * z * (2 * zz - 3 * xx - 3 * yy), _C3[4] * x * (4 * zz - xx - yy), _C3[5] * z * (xx - yy),
                     _C3[6] * x * (xx - 3 * yy)], axis=1)


def _factor_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        orc.set_num_threads(2)
        scene, cams = syn.make_config(4, P=1500, W=96, H=64, n_cameras=N_VIEWS)
        factors = []
        for v in D.views_for_rank(N_VIEWS, rank, world):
            g, _ = _view_grads(scene, cams[v], 100 + v)
            factors.append((g["shs"][:, 0, :] / _C0, cams[v].campos, 3))  # band 0 of dL_dsh is C0 * the factor
        g_all, c_all, degree = D.gather_sh_factors(factors, 1500, torch.device("cpu"))
        torch.save(dict(g=g_all, c=c_all, degree=degree), os.path.join(out_dir, f"f{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_gathered_sh_factors_rebuild_the_summed_sh_gradient(tmp_path):
    world = 2
    mp.spawn(_factor_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    f0, f1 = torch.load(tmp_path / "f0.pt"), torch.load(tmp_path / "f1.pt")
    assert torch.equal(f0["g"], f1["g"]) and torch.equal(f0["c"], f1["c"]) and f0["degree"] == 3
    assert f0["g"].shape == (N_VIEWS, 1500, 3) and f0["c"].shape == (N_VIEWS, 3)
    scene, cams = syn.make_config(4, P=1500, W=96, H=64, n_cameras=N_VIEWS)
    order = [v for r in range(world) for v in D.views_for_rank(N_VIEWS, r, world)]  # rank-major, as all_gather lays out
    for i, v in enumerate(order):
        assert torch.equal(f0["c"][i], cams[v].campos.reshape(3).float())
    want = np.zeros((1500, 16, 3))
    got = np.zeros((1500, 16, 3))
    means = scene["means3D"].numpy().astype(np.float64)
    for i, v in enumerate(order):
        g, _ = _view_grads(scene, cams[v], 100 + v)
        want += g["shs"].numpy()
        d = means - f0["c"][i].numpy().astype(np.float64)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        got += _sh_basis(d)[:, :, None] * f0["g"][i].numpy().astype(np.float64)[:, None, :]
    assert np.abs(want).max() > 0
    assert np.abs(got - want).max() <= 2e-5 * np.abs(want).max()


def test_factor_sink_is_scoped():
    from langsplat_b200 import rasterizer as RZ
    assert RZ._sh_factor_sink is None
    with D.factored_sh_gradients() as outer:
        assert RZ._sh_factor_sink is outer
        with D.factored_sh_gradients() as inner:
            assert RZ._sh_factor_sink is inner
        assert RZ._sh_factor_sink is outer
    assert RZ._sh_factor_sink is None
    g, c, deg = D.gather_sh_factors([], 10, torch.device("cpu"))
    assert g.shape == (0, 10, 3) and c.shape == (0, 3) and deg == 0
