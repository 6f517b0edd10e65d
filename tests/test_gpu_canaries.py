"""GPU: out-of-bounds WRITE check with guard bands (compute-sanitizer is closed on this pool).

Every buffer the host layer hands to liblsr.so — outputs, geom / image / list state, the scratch workspace, the packed
gradient allocation — is allocated with a 4 KB band of a known byte pattern on either side; after forward and backward
every band must be untouched.  Covers the configurations whose buffer sets differ: RGB + feature, RGB only (the
reference's own out-of-bounds case: backward.cu:494-495 reads the feature gradient although include_feature is off),
feature-only training, precomputed colour / covariance, a ragged image, an empty scene, the asynchronous forward with
a list that is too small (stores must be clipped at the capacity), several feature levels.
"""
import pytest
import torch

import langsplat_b200 as lsb
from helpers import R, run_module, settings_for, syn

pytestmark = pytest.mark.gpu

GUARD = 4096  # bytes on each side; a multiple of 256 so that the inner buffer keeps the allocator's alignment
PATTERN = 0xA5


class _Guarded:
    """Stand-in for torch.empty / torch.zeros inside langsplat_b200.rasterizer: same call, guarded storage."""

    def __init__(self):
        self.bands = []
        self._empty, self._zeros = torch.empty, torch.zeros

    def _alloc(self, fill, *size, **kw):
        if len(size) == 1 and isinstance(size[0], (tuple, list, torch.Size)):
            size = tuple(size[0])
        dtype = kw.get("dtype", torch.float32)
        device = kw.get("device", None)
        if device is None or torch.device(device).type != "cuda":
            return (self._zeros if fill else self._empty)(*size, **kw)
        n = 1
        for d in size:
            n *= int(d)
        esize = torch.empty((), dtype=dtype).element_size()
        inner = (n * esize + 255) // 256 * 256
        raw = self._empty((GUARD + inner + GUARD,), dtype=torch.uint8, device=device)
        raw[:GUARD] = PATTERN
        raw[GUARD + n * esize:] = PATTERN
        self.bands.append((raw, n * esize))
        t = raw[GUARD:GUARD + n * esize].view(dtype).view(*size) if n else self._empty(*size, **kw)
        if fill and n:
            t.zero_()
        return t

    def empty(self, *size, **kw):
        return self._alloc(False, *size, **kw)

    def zeros(self, *size, **kw):
        return self._alloc(True, *size, **kw)

    def check(self):
        torch.cuda.synchronize()
        assert self.bands, "nothing was allocated through the guarded allocator"
        for raw, nbytes in self.bands:
            assert bool((raw[:GUARD] == PATTERN).all()), f"write below a {nbytes}-byte buffer"
            assert bool((raw[GUARD + nbytes:] == PATTERN).all()), f"write beyond a {nbytes}-byte buffer"


@pytest.fixture
def guarded(monkeypatch):
    g = _Guarded()

    class _TorchProxy:
        def __getattr__(self, name):
            if name == "empty":
                return g.empty
            if name == "zeros":
                return g.zeros
            return getattr(torch, name)

    monkeypatch.setattr(R, "torch", _TorchProxy())
    monkeypatch.setattr(R, "_scratch", {})  # the cached workspace must come from the guarded allocator too
    yield g
    lsb.set_async_forward(False, 1.5)
    for pend in R._pending.values():
        pend.clear()


def _grads(cam, seed, feature_only=False):
    return syn.make_grads(cam.image_width, cam.image_height, seed, feature_only=feature_only)


@pytest.mark.parametrize("include_feature", [True, False])
@pytest.mark.parametrize("W,H", [(256, 256), (250, 130)])
def test_forward_backward_stay_inside_their_buffers(guarded, include_feature, W, H):
    scene, cams = syn.make_config(1, W=W, H=H)
    dcol, dfeat = _grads(cams[0], 3)
    requires = ("means3D", "shs", "opacities", "scales", "rotations") + (("language_feature",) if include_feature else ())
    out = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cams[0], grads=(dcol, dfeat),
                     include_feature=include_feature, requires=requires, bg=(0.2, 0.3, 0.4))
    assert torch.isfinite(out["color"]).all()
    guarded.check()


def test_feature_only_and_precomputed_branches(guarded):
    scene, cams = syn.make_config(3, P=30_000, W=360, H=270)
    _, dfeat = _grads(cams[0], 4, feature_only=True)
    run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cams[0], grads=(None, dfeat),
               requires=("language_feature",))
    guarded.check()
    P = scene["means3D"].shape[0]
    g = torch.Generator().manual_seed(1)
    cp = torch.rand(P, 3, generator=g)
    L = torch.randn(P, 3, 3, generator=g) * 0.02
    cov = L @ L.transpose(1, 2)
    c3 = torch.stack([cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2]], 1).contiguous()
    dcol, dfeat = _grads(cams[0], 5)
    run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cams[0], use_colors_precomp=cp,
               use_cov3D_precomp=c3, grads=(dcol, dfeat),
               requires=("means3D", "colors_precomp", "cov3D_precomp", "opacities", "language_feature"))
    guarded.check()


def test_empty_scene_and_feature_levels(guarded):
    scene, cams = syn.make_config(1, P=4000, W=128, H=96)
    dev = torch.device("cuda")
    rast = R.GaussianRasterizer(settings_for(cams[0], dev))
    d = {k: v.to(dev) for k, v in scene.items()}
    levels = [d["language_feature"], d["language_feature"].flip(0).contiguous(), -d["language_feature"]]
    rast.forward_feature_levels(means3D=d["means3D"], opacities=d["opacities"], language_features=levels, shs=d["shs"],
                                scales=d["scales"], rotations=d["rotations"])
    empty = {k: v[:0].to(dev) for k, v in scene.items()}
    color, feat, radii = rast(means3D=empty["means3D"], means2D=torch.zeros(0, 3, device=dev), shs=empty["shs"],
                              language_feature_precomp=empty["language_feature"], opacities=empty["opacities"],
                              scales=empty["scales"], rotations=empty["rotations"])
    assert float(color.abs().max()) == 0.0
    guarded.check()


def test_async_forward_clips_an_overflowing_list(guarded):
    """The asynchronous forward sizes its list from earlier frames; when the frame turns out larger, the duplication
    kernel must clip its stores at the capacity and the sort must stay inside the clipped list."""
    scene, cams = syn.make_config(1)
    run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cams[0])  # synchronous first frame
    R._recent_R[torch.cuda.current_device()].clear()
    lsb.reserve_instances("cuda", 1000)  # -> capacity 64 Ki, the frame has ~84 k instances
    with lsb.async_forward(True, 1.0):
        out = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cams[0])
        assert torch.isnan(out["color"]).all()
        with pytest.raises(lsb.LsrError):
            lsb.check_async_status()
        dcol, dfeat = _grads(cams[0], 6)
        ok = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cams[0], grads=(dcol, dfeat))
        lsb.check_async_status()
        assert torch.isfinite(ok["color"]).all()
    guarded.check()


@pytest.mark.parametrize("P", [8001, 4099])
def test_odd_gaussian_counts_keep_every_vector_access_aligned(guarded, P):
    """The backward's buffers are slices of one allocation whose offsets depend on P: with an odd P the float2 / float4
    accesses (dL_dcov3D, dL_drotations, the accumulation record) must still land on aligned addresses — dense SH
    gradient, factored SH gradient (no colour buffer at all) and precomputed colours."""
    from langsplat_b200 import dist as D
    scene, cams = syn.make_config(4, P=P, W=160, H=120, n_cameras=1)
    dcol, dfeat = _grads(cams[0], 8)
    names = ("means3D", "shs", "opacities", "scales", "rotations", "language_feature")
    dense = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cams[0], grads=(dcol, dfeat),
                       requires=names)
    with D.factored_sh_gradients() as factors:
        fact = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cams[0], grads=(dcol, dfeat),
                          requires=names)
    assert len(factors) == 1
    for k in ("means3D", "scales", "rotations", "opacities", "language_feature"):
        assert torch.isfinite(fact["grads"][k]).all() and torch.isfinite(dense["grads"][k]).all()
    cp = torch.rand(P, 3, generator=torch.Generator().manual_seed(2))
    run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cams[0], use_colors_precomp=cp,
               grads=(dcol, dfeat), requires=("means3D", "colors_precomp", "opacities", "scales", "rotations"))
    guarded.check()


@pytest.mark.parametrize("mode", ["feature_only", "means2D_and_colors"])
def test_accumulation_record_without_the_per_gaussian_kernel(guarded, monkeypatch, mode):
    """lsr_grads::blend_acc when no per-Gaussian kernel follows the blend kernel (feature-only training, or only the
    screen-space / precomputed-colour gradients wanted): unpack_acc_kernel writes the tensors.  Same numbers as the
    five-accumulator path, up to the order of the float atomics."""
    scene, cams = syn.make_config(3, P=30_011, W=360, H=270)
    kw = {}
    if mode == "feature_only":
        grads, requires = (None, _grads(cams[0], 4, feature_only=True)[1]), ("language_feature",)
    else:
        grads, requires = _grads(cams[0], 4), ("colors_precomp", "language_feature")
        kw["use_colors_precomp"] = torch.rand(30_011, 3, generator=torch.Generator().manual_seed(1))
    out = {}
    for setting in (0, 2, 0):
        monkeypatch.setattr(R, "_USE_BLEND_ACC", setting)
        out.setdefault(setting, []).append(run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene,
                                                      cams[0], grads=grads, requires=requires, **kw)["grads"])
    for k in requires + ("means2D",):
        a, b, again = out[0][0][k], out[2][0][k], out[0][1][k]
        assert b is not None and torch.isfinite(b).all()
        noise = float((a - again).abs().max()) / max(float(a.abs().max()), 1e-30)
        assert float((a - b).abs().max()) <= max(1e-5, 4.0 * noise) * float(a.abs().max()), k
    guarded.check()
