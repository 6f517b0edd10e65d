"""Ground-truth language-feature map: the oracle's restatement against golden vectors produced by
the reference's own Camera.get_language_feature (tests/golden/make_golden_gt_features.py), then — on the GPU —
lsr_language_feature_map through langsplat_b200.gt_features against the same vectors (bit-exact: it is a gather),
the per-camera cache, the drop-in `get_language_feature`, and PyTorch's error behaviour for out-of-range indices."""
import os
import types

import numpy as np
import pytest
import torch

from helpers import orc

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "gt_features.npz"))
CASES = ("full_37x53", "crop_40x64", "int_24x31")


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_vectors(name):
    H, W = (int(v) for v in GOLD[f"{name}.hw"])
    for level in range(4):
        feat, mask = orc.language_feature_map(GOLD[f"{name}.seg"], GOLD[f"{name}.table"], level, H, W)
        assert np.array_equal(feat, GOLD[f"{name}.l{level}.feature"])
        assert np.array_equal(mask, GOLD[f"{name}.l{level}.mask"])
    with pytest.raises(ValueError):
        orc.language_feature_map(GOLD[f"{name}.seg"], GOLD[f"{name}.table"], 4, H, W)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_gpu_matches_reference_vectors(name):
    from langsplat_b200.gt_features import language_feature_map
    dev = torch.device("cuda:0")
    H, W = (int(v) for v in GOLD[f"{name}.hw"])
    seg, table = torch.from_numpy(GOLD[f"{name}.seg"]).to(dev), torch.from_numpy(GOLD[f"{name}.table"]).to(dev)
    for level in range(4):
        feat, mask, mask_f = language_feature_map(seg, table, level, H, W, float_mask=True)
        assert feat.shape == (3, H, W) and mask.shape == (1, H, W) and mask.dtype == torch.bool
        assert np.array_equal(feat.cpu().numpy(), GOLD[f"{name}.l{level}.feature"])
        assert np.array_equal(mask.cpu().numpy(), GOLD[f"{name}.l{level}.mask"])
        assert torch.equal(mask_f, mask.float())
    with pytest.raises(ValueError):
        language_feature_map(seg, table, 4, H, W)


@pytest.mark.gpu
def test_gpu_cache_dropin_and_errors(tmp_path):
    from langsplat_b200 import LsrError
    from langsplat_b200.gt_features import LanguageFeatureCache, get_language_feature, language_feature_map
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(3)
    K, H, W = 301, 270, 480
    seg = rng.integers(-1, K, size=(4, H, W)).astype(np.float32)
    table = rng.standard_normal((K, 3)).astype(np.float32)
    np.save(tmp_path / "frame_00001_s.npy", seg)
    np.save(tmp_path / "frame_00001_f.npy", table)
    cam = types.SimpleNamespace(image_name="frame_00001", image_height=H, image_width=W)
    cache = LanguageFeatureCache(dev)
    feat, mask = get_language_feature(cam, str(tmp_path), 1, cache=cache)
    want_f, want_m = orc.language_feature_map(seg, table, 1, H, W)
    assert np.array_equal(feat.cpu().numpy(), want_f) and np.array_equal(mask.cpu().numpy(), want_m)
    os.remove(tmp_path / "frame_00001_s.npy")  # second request: nothing is reloaded, the same tensors come back
    feat2, mask2 = get_language_feature(cam, str(tmp_path), 1, cache=cache)
    assert feat2 is feat and mask2 is mask
    feat3, _ = get_language_feature(cam, str(tmp_path), 3, cache=cache)  # other level: inputs are resident
    assert np.array_equal(feat3.cpu().numpy(), orc.language_feature_map(seg, table, 3, H, W)[0])
    # PyTorch raises for indices outside [-K, K); so does the validated call, the unvalidated one writes zeros
    bad = torch.from_numpy(seg).to(dev)
    bad[0, 5, 7] = K
    bad[0, 6, 8] = -K - 1
    with pytest.raises(IndexError):
        language_feature_map(bad, torch.from_numpy(table).to(dev), 0)
    f0, m0 = language_feature_map(bad, torch.from_numpy(table).to(dev), 0, validate=False)
    assert float(f0[:, 5, 7].abs().max()) == 0.0 and bool(m0[0, 5, 7])
    neg = torch.from_numpy(seg).to(dev)
    neg[0, 0, 0] = -2  # a legal negative index: second-to-last row, mask stays True (only -1 is "no mask")
    f1, m1 = language_feature_map(neg, torch.from_numpy(table).to(dev), 0)
    assert torch.equal(f1[:, 0, 0].cpu(), torch.from_numpy(table[K - 2])) and bool(m1[0, 0, 0])
    with pytest.raises(IndexError):
        language_feature_map(torch.from_numpy(seg).to(dev), torch.from_numpy(table).to(dev), 0, H + 1, W)
    with pytest.raises(LsrError):
        language_feature_map(torch.from_numpy(seg), torch.from_numpy(table), 0)  # CPU tensors: no CPU path
