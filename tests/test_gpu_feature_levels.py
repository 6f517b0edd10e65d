"""GPU: several language-feature levels on shared geometry (`GaussianRasterizer.forward_feature_levels` ->
lsr_forward_feature_level).  Every level's feature image must equal what a separate full forward
with that level's tensor returns (same kernels, same decisions: bit-identical), and match the CPU oracle to the image
tolerance (1e-5).  Covers a ragged image size, a background colour and the error paths."""
import numpy as np
import pytest
import torch

from helpers import R, oracle_forward, settings_for, syn

pytestmark = pytest.mark.gpu


def _levels(P, n, seed):
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n):
        f = torch.randn(P, 3, generator=g)
        out.append(f / f.norm(dim=1, keepdim=True))
    return out


@pytest.mark.parametrize("P,W,H,bg", [(16384, 256, 256, (0.0, 0.0, 0.0)), (5000, 203, 131, (0.3, 0.1, 0.7))])
def test_levels_equal_separate_forwards_and_oracle(P, W, H, bg):
    dev = torch.device("cuda:0")
    scene, cams = syn.make_config(1, P=P, W=W, H=H)
    cam = cams[0]
    levels = _levels(P, 3, 99)
    rs = settings_for(cam, dev, bg=bg)
    d = {k: v.to(dev) for k, v in scene.items()}
    rast = R.GaussianRasterizer(rs)
    color, feats, radii = rast.forward_feature_levels(
        means3D=d["means3D"], opacities=d["opacities"], language_features=[f.to(dev) for f in levels], shs=d["shs"],
        scales=d["scales"], rotations=d["rotations"])
    assert len(feats) == 3
    for i, lang in enumerate(levels):
        with torch.no_grad():
            c2, f2, r2 = rast(means3D=d["means3D"], means2D=torch.zeros_like(d["means3D"]), shs=d["shs"],
                              language_feature_precomp=lang.to(dev), opacities=d["opacities"], scales=d["scales"],
                              rotations=d["rotations"])
        assert torch.equal(feats[i], f2), f"level {i} differs from a separate forward"
        assert torch.equal(color, c2) and torch.equal(radii, r2)
        fw = oracle_forward(dict(scene, language_feature=lang), cam, bg=bg)
        np.testing.assert_allclose(feats[i].cpu().numpy(), fw["feature"], atol=1e-5, rtol=0)


def test_levels_all_culled_and_errors():
    dev = torch.device("cuda:0")
    scene, cams = syn.make_config(1, P=512, W=64, H=64)
    cam = cams[0]
    d = {k: v.to(dev) for k, v in scene.items()}
    behind = d["means3D"] + 1000.0 * (cam.campos.to(dev) / cam.campos.norm())  # everything behind the camera
    rast = R.GaussianRasterizer(settings_for(cam, dev))
    color, feats, radii = rast.forward_feature_levels(
        means3D=behind, opacities=d["opacities"], language_features=[d["language_feature"]] * 2, shs=d["shs"],
        scales=d["scales"], rotations=d["rotations"])
    assert int((radii > 0).sum()) == 0 and all(float(f.abs().max()) == 0.0 for f in feats)
    with pytest.raises(R.LsrError):
        rast.forward_feature_levels(means3D=d["means3D"], opacities=d["opacities"], language_features=[],
                                    shs=d["shs"], scales=d["scales"], rotations=d["rotations"])
    with pytest.raises(R.LsrError):
        rast.forward_feature_levels(means3D=d["means3D"], opacities=d["opacities"],
                                    language_features=[d["language_feature"], d["language_feature"][:10]],
                                    shs=d["shs"], scales=d["scales"], rotations=d["rotations"])
    with pytest.raises(R.LsrError):
        R.GaussianRasterizer(settings_for(cam, dev, include_feature=False)).forward_feature_levels(
            means3D=d["means3D"], opacities=d["opacities"], language_features=[d["language_feature"]], shs=d["shs"],
            scales=d["scales"], rotations=d["rotations"])
