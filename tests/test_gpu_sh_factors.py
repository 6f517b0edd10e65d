"""GPU: factored SH gradients (langsplat_b200.dist.factored_sh_gradients / finish_sh_gradients ->
lsr_grads.dL_dcolors_masked + lsr_sh_gradient_from_views).  One view's dL_dsh is basis(dir) (x) g per Gaussian
(RAST cuda_rasterizer/backward.cu:31-129 restated in oracle/oracle.py), so the sum over the views of a step formed from
the 12-byte factors must equal the sum of the dense per-view gradients up to float summation order, every other
gradient must be unchanged by the factoring, and the oracle's SH gradient pins both."""
import numpy as np
import pytest
import torch

from helpers import R, oracle_forward, orc, rel_err, settings_for, syn
from langsplat_b200 import dist as D

pytestmark = pytest.mark.gpu

NAMES = ("means3D", "shs", "opacities", "scales", "rotations", "language_feature")


def _step(scene, cams, dev, degree, factored):
    params = {k: scene[k].to(dev).requires_grad_(True) for k in NAMES}
    means2D = torch.zeros_like(params["means3D"], requires_grad=True)

    def run():
        for i, cam in enumerate(cams):
            dcol, dfeat = [t.to(dev) for t in syn.make_grads(cam.image_width, cam.image_height, 300 + i)]
            rast = R.GaussianRasterizer(settings_for(cam, dev, sh_degree=degree))
            color, feat, radii = rast(means3D=params["means3D"], means2D=means2D, shs=params["shs"],
                                      language_feature_precomp=params["language_feature"],
                                      opacities=params["opacities"], scales=params["scales"],
                                      rotations=params["rotations"])
            torch.autograd.backward([color, feat], [dcol, dfeat])

    if factored:
        with D.factored_sh_gradients() as factors:
            run()
        assert params["shs"].grad is None and len(factors) == len(cams)
        params["shs"].grad = D.finish_sh_gradients(factors, params["means3D"], params["shs"].shape[1])
    else:
        run()
    return {k: p.grad.detach().cpu() for k, p in params.items()}


@pytest.mark.parametrize("degree", [0, 1, 2, 3])
def test_factored_step_equals_dense_step(degree):
    dev = torch.device("cuda:0")
    scene, cams = syn.make_config(4, P=20000, W=200, H=136, n_cameras=3)
    dense = _step(scene, cams, dev, degree, factored=False)
    again = _step(scene, cams, dev, degree, factored=False)
    fact = _step(scene, cams, dev, degree, factored=True)
    # same kernels, same inputs; the blend backward sums with unordered atomics, so two dense runs already differ in
    # the last bits: the bar is that run-to-run noise (floored at 1e-5), far below what a missing term would show
    for k in NAMES:
        if k != "shs":
            tol = max(1e-5, 4.0 * rel_err(again[k], dense[k]))
            assert rel_err(fact[k], dense[k]) <= tol, f"{k} changed under SH factoring"
    assert fact["shs"].shape == dense["shs"].shape
    n = (degree + 1) ** 2
    assert float(fact["shs"][:, n:].abs().max()) == 0.0 if n < 16 else True  # unused bands: exact zeros, like dense
    assert rel_err(fact["shs"], dense["shs"]) <= max(1e-5, 4.0 * rel_err(again["shs"], dense["shs"]))


def test_single_view_factor_matches_oracle():
    dev = torch.device("cuda:0")
    scene, cams = syn.make_config(1, P=6000, W=128, H=96)
    cam = cams[0]
    got = _step(scene, [cam], dev, 3, factored=True)
    dcol, dfeat = syn.make_grads(cam.image_width, cam.image_height, 300)
    fw = oracle_forward(scene, cam)
    g = orc.backward(fw, dcol.numpy(), dfeat.numpy())
    want = np.asarray(g["shs"], np.float32).reshape(scene["shs"].shape)
    assert rel_err(got["shs"], want) <= 1e-4  # the gradient bar of the path


def test_no_views_and_bad_arguments():
    dev = torch.device("cuda:0")
    means = torch.randn(100, 3, device=dev)
    assert D.finish_sh_gradients([], means, 16) is None
    from langsplat_b200 import _lib as L
    import ctypes
    lib = L.load()
    out = torch.empty(100, 4, 3, device=dev)
    rc = lib.lsr_sh_gradient_from_views(100, 4, 3, 0, ctypes.c_void_p(means.data_ptr()), None, None,
                                        ctypes.c_void_p(out.data_ptr()), None)
    assert rc != L.LSR_OK and b"SH coefficients" in L.last_error().encode()
    out = torch.full((100, 16, 3), 7.0, device=dev)  # zero views: the sum is zero
    rc = lib.lsr_sh_gradient_from_views(100, 16, 3, 0, ctypes.c_void_p(means.data_ptr()), None, None,
                                        ctypes.c_void_p(out.data_ptr()),
                                        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == L.LSR_OK
    torch.cuda.synchronize()
    assert float(out.abs().max()) == 0.0


# ---- per-view pointers and the peer buffer / barrier (single process; the 2-rank run is tests/peer_exchange_check.py,
# launched under torchrun by profiles/scale_ab.sh on a multi-GPU box) ----
def _ptr_array(ptrs):
    import ctypes
    return (ctypes.c_void_p * len(ptrs))(*ptrs)


def test_view_pointers_equal_contiguous_and_chunking():
    import ctypes
    from langsplat_b200 import _lib as L
    lib = L.load()
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(5)
    P, V = 3000, 70  # more than LSR_MAX_VIEWS: second chunk accumulates
    means = torch.randn(P, 3, generator=g).to(dev)
    cams = (4.0 * torch.randn(V, 3, generator=g)).to(dev)
    fac = torch.randn(V, P, 3, generator=g)
    fac[torch.rand(V, P, generator=g) < 0.4] = 0.0  # culled in that view
    fac = fac.to(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def contiguous(v0, v1, M=16, D=3):
        out = torch.empty(P, M, 3, device=dev)
        rc = lib.lsr_sh_gradient_from_views(P, M, D, v1 - v0, ctypes.c_void_p(means.data_ptr()),
                                            ctypes.c_void_p(cams[v0:v1].contiguous().data_ptr()),
                                            ctypes.c_void_p(fac[v0:v1].contiguous().data_ptr()),
                                            ctypes.c_void_p(out.data_ptr()), stream)
        assert rc == L.LSR_OK, L.last_error()
        return out

    def by_pointers(views, M=16, D=3):
        out = torch.empty(P, M, 3, device=dev)
        gp = _ptr_array([fac[v].data_ptr() for v in views])
        cp = _ptr_array([cams[v].data_ptr() for v in views])
        rc = lib.lsr_sh_gradient_from_view_ptrs(P, M, D, len(views), ctypes.c_void_p(means.data_ptr()), gp, cp,
                                                ctypes.c_void_p(out.data_ptr()), stream)
        assert rc == L.LSR_OK, L.last_error()
        return out

    assert torch.equal(by_pointers(list(range(8))), contiguous(0, 8))  # same kernel, same order
    assert torch.equal(by_pointers(list(range(8)), M=9, D=2), contiguous(0, 8, M=9, D=2))
    whole = by_pointers(list(range(V)))
    parts = contiguous(0, 64) + contiguous(64, V)
    assert rel_err(whole, parts) <= 1e-6
    # float64 restatement of basis(dir) (x) g for a handful of Gaussians
    from test_dist_gloo import _sh_basis
    idx = torch.arange(0, P, 97)
    want = np.zeros((len(idx), 16, 3))
    for v in range(V):
        d = (means[idx] - cams[v]).double().cpu().numpy()
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        want += _sh_basis(d)[:, :, None] * fac[v, idx].double().cpu().numpy()[:, None, :]
    assert rel_err(whole[idx], want) <= 1e-5
    bad = _ptr_array([fac[0].data_ptr(), 0])
    rc = lib.lsr_sh_gradient_from_view_ptrs(P, 16, 3, 2, ctypes.c_void_p(means.data_ptr()), bad, bad,
                                            ctypes.c_void_p(whole.data_ptr()), stream)
    assert rc != L.LSR_OK and "view 1" in L.last_error()


def test_peer_barrier_two_ranks_on_two_streams_and_timeout():
    import ctypes
    from langsplat_b200 import _lib as L
    lib = L.load()
    bufs, hnd = [], (ctypes.c_uint8 * L.PEER_HANDLE_BYTES)()
    for _ in range(2):
        p = ctypes.c_void_p()
        assert lib.lsr_peer_alloc(4096, ctypes.byref(p), hnd) == L.LSR_OK, L.last_error()
        bufs.append(p.value)
    assert any(hnd)  # a real IPC handle came back
    headers = _ptr_array(bufs)
    s0, s1 = torch.cuda.Stream(), torch.cuda.Stream()
    for epoch in (1, 2, 3):  # both "ranks" arrive (two streams of one device): the barrier completes
        assert lib.lsr_peer_barrier(0, 2, headers, epoch, ctypes.c_void_p(s0.cuda_stream)) == L.LSR_OK
        assert lib.lsr_peer_barrier(1, 2, headers, epoch, ctypes.c_void_p(s1.cuda_stream)) == L.LSR_OK
    torch.cuda.synchronize()
    for r in (0, 1):
        assert lib.lsr_peer_status(ctypes.c_void_p(bufs[r]), None) == L.LSR_OK
    # rank 1 never arrives at epoch 4: rank 0 gives up after the time-out instead of hanging the device
    assert lib.lsr_peer_barrier(0, 2, headers, 4, ctypes.c_void_p(s0.cuda_stream)) == L.LSR_OK
    torch.cuda.synchronize()
    assert lib.lsr_peer_status(ctypes.c_void_p(bufs[0]), None) != L.LSR_OK and "timed out" in L.last_error()
    assert lib.lsr_peer_barrier(0, 17, headers, 5, None) != L.LSR_OK
    assert lib.lsr_peer_barrier(2, 2, headers, 5, None) != L.LSR_OK
    for b in bufs:
        assert lib.lsr_peer_free(ctypes.c_void_p(b)) == L.LSR_OK


def test_peer_exchange_single_rank_equals_list_sink():
    dev = torch.device("cuda:0")
    scene, cams = syn.make_config(4, P=8001, W=160, H=120, n_cameras=2)  # odd P: unaligned second slot
    listed = _step(scene, cams, dev, 3, factored=True)
    ex = D.PeerFactorExchange(8001, 2, dev)
    try:
        for _ in range(3):  # slots alternate; every step must give the same answer
            params = {k: scene[k].to(dev).requires_grad_(True) for k in NAMES}
            means2D = torch.zeros_like(params["means3D"], requires_grad=True)
            with D.factored_sh_gradients(ex.begin_step()):
                for i, cam in enumerate(cams):
                    dcol, dfeat = [t.to(dev) for t in syn.make_grads(cam.image_width, cam.image_height, 300 + i)]
                    rast = R.GaussianRasterizer(settings_for(cam, dev, sh_degree=3))
                    color, feat, radii = rast(means3D=params["means3D"], means2D=means2D, shs=params["shs"],
                                              language_feature_precomp=params["language_feature"],
                                              opacities=params["opacities"], scales=params["scales"],
                                              rotations=params["rotations"])
                    torch.autograd.backward([color, feat], [dcol, dfeat])
            got = ex.finish(params["means3D"], 16)
            assert rel_err(got, listed["shs"]) <= 1e-5
        ex.check()
        with pytest.raises(R.LsrError):
            ex.begin_step().finish(params["means3D"], 16)  # no backward ran
    finally:
        ex.close()


def test_peer_allreduce_two_ranks_in_one_process():
    """lsr_peer_allreduce with both 'ranks' on one device: each call reduces its own chunk from both inputs and stores
    it into both outputs; together they leave the full sum everywhere (odd length in float4s, tail chunk shorter)."""
    import ctypes
    from langsplat_b200 import _lib as L
    lib = L.load()
    dev = torch.device("cuda:0")
    n = 4 * 100003
    g = torch.Generator().manual_seed(3)
    ins = [torch.randn(n, generator=g).to(dev) for _ in range(2)]
    outs = [torch.full((n,), float("nan"), device=dev) for _ in range(2)]
    ip, op = _ptr_array([t.data_ptr() for t in ins]), _ptr_array([t.data_ptr() for t in outs])
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for rank in (0, 1):
        assert lib.lsr_peer_allreduce(rank, 2, ip, op, n, stream) == L.LSR_OK, L.last_error()
    torch.cuda.synchronize()
    want = ins[0] + ins[1]
    assert torch.equal(outs[0], want) and torch.equal(outs[1], want)
    # generic world size (not one of the unrolled 2 / 4 / 8): three ranks
    ins3 = ins + [torch.randn(n, generator=g).to(dev)]
    outs3 = [torch.zeros(n, device=dev) for _ in range(3)]
    ip3, op3 = _ptr_array([t.data_ptr() for t in ins3]), _ptr_array([t.data_ptr() for t in outs3])
    for rank in range(3):
        assert lib.lsr_peer_allreduce(rank, 3, ip3, op3, n, stream) == L.LSR_OK
    torch.cuda.synchronize()
    want3 = (ins3[0] + ins3[1]) + ins3[2]
    assert all(torch.equal(o, want3) for o in outs3)
    assert lib.lsr_peer_allreduce(0, 2, ip, op, n + 2, stream) != L.LSR_OK  # not a multiple of 4
    bad = _ptr_array([ins[0].data_ptr() + 4, ins[1].data_ptr()])
    assert lib.lsr_peer_allreduce(0, 2, bad, op, n, stream) != L.LSR_OK and "aligned" in L.last_error()
    assert lib.lsr_peer_allreduce(0, 1, ip, op, 0, stream) == L.LSR_OK  # nothing to do


def test_exchange_with_device_epochs_replays_from_cuda_graphs():
    """PeerFactorExchange(device_epochs=True): the barriers count their epochs on the device, the payload slot is
    chosen by the caller, so forward + backward + exchange of a step is captured once per slot and replayed — here with
    one rank (the barrier signals and waits on itself; the multi-rank case runs in bench.py --gpus N)."""
    from langsplat_b200.graphs import GraphedStep
    dev = torch.device("cuda:0")
    P = 8001
    scene, cams = syn.make_config(4, P=P, W=160, H=120, n_cameras=2)
    listed = _step(scene, cams, dev, 3, factored=True)
    others = tuple(k for k in NAMES if k != "shs")
    ex = D.PeerFactorExchange(P, 2, dev, device_epochs=True)
    try:
        params = {k: scene[k].to(dev).requires_grad_(True) for k in NAMES}
        means2D = torch.zeros_like(params["means3D"], requires_grad=True)
        grads_in = [[t.to(dev) for t in syn.make_grads(c.image_width, c.image_height, 300 + i)]
                    for i, c in enumerate(cams)]
        rasts = [R.GaussianRasterizer(settings_for(c, dev, sh_degree=3)) for c in cams]

        def make(slot):
            def fn():
                acc = None
                with D.factored_sh_gradients(ex.begin_step(slot=slot)):
                    for rast, (dcol, dfeat) in zip(rasts, grads_in):
                        color, feat, radii = rast(means3D=params["means3D"], means2D=means2D, shs=params["shs"],
                                                  language_feature_precomp=params["language_feature"],
                                                  opacities=params["opacities"], scales=params["scales"],
                                                  rotations=params["rotations"])
                        g = torch.autograd.grad([color, feat], [params[k] for k in others], [dcol, dfeat])
                        acc = list(g) if acc is None else [a + b for a, b in zip(acc, g)]
                sh, _ = ex.exchange(params["means3D"], 16, [])
                return sh, acc
            return fn

        steps = [GraphedStep(make(0)), GraphedStep(make(1))]
        for i in (0, 1, 0, 1, 1):
            sh, acc = steps[i].replay()
            steps[i].check()
            assert rel_err(sh.cpu(), listed["shs"]) <= 1e-5
            for k, g in zip(others, acc):
                assert rel_err(g.cpu(), listed[k]) <= 1e-4, k
        ex.check()
    finally:
        ex.close()
