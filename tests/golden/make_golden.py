#!/usr/bin/env python
"""Generates the golden vectors under tests/golden/ by running the UNMODIFIED reference CUDA extension
(baseline/_ref, built from /root/reference/submodules/langsplat-rasterization by baseline/build_ref.sh) on a B200.

The reference ships no tests, golden vectors or CPU implementation of this path, so its own outputs on
seeded inputs are the pin: each fixture stores the inputs and everything the reference produced for them — radii, the
parsed geometry / binning / image buffers (depths, means2D, cov3D, conic, rgb, tiles_touched, sorted keys, point list,
tile ranges, final_T, n_contrib), both images and all nine gradients.

Run on the GPU box (the reference needs a GPU):
    gpurun -- 'python tests/golden/make_golden.py gpurun_out/golden'
then copy gpurun_out/golden/*.npz into tests/golden/.  tests/test_oracle_golden.py (CPU) checks the oracle against
them; tests/test_gpu_golden.py checks the CUDA path.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from helpers import load_reference, ref_forward_buffers, run_module, syn  # noqa: E402

# name -> (scene seed config, P, W, H, sh_degree, include_feature, bg, feature_only, grad seed)
CASES = {
    "rgb_feat_bg": dict(cfg=1, P=1024, W=64, H=64, deg=3, feat=True, bg=(0.1, 0.4, 0.7), feature_only=False, gs=31),
    "feature_only_ragged": dict(cfg=3, P=900, W=72, H=40, deg=3, feat=True, bg=(0.0, 0.0, 0.0), feature_only=True,
                                gs=32),
    "rgb_only_deg1": dict(cfg=2, P=800, W=48, H=48, deg=1, feat=False, bg=(0.0, 0.0, 0.0), feature_only=False, gs=33),
}


def main(out_dir: str) -> None:
    REF = load_reference()
    assert REF is not None, "baseline/_ref is not built"
    os.makedirs(out_dir, exist_ok=True)
    for name, c in CASES.items():
        scene, cams = syn.make_config(c["cfg"], P=c["P"], W=c["W"], H=c["H"])
        cam = cams[0]
        dcol, dfeat = syn.make_grads(c["W"], c["H"], c["gs"], feature_only=c["feature_only"])
        buf = ref_forward_buffers(REF, scene, cam, include_feature=c["feat"], sh_degree=c["deg"], bg=c["bg"])
        requires = ("language_feature",) if c["feature_only"] else (
            "means3D", "shs", "language_feature", "opacities", "scales", "rotations")
        if not c["feat"]:
            requires = tuple(k for k in requires if k != "language_feature")
        out = run_module(REF.GaussianRasterizer, REF.GaussianRasterizationSettings, scene, cam,
                         include_feature=c["feat"], sh_degree=c["deg"], bg=c["bg"],
                         grads=(None if c["feature_only"] else dcol, dfeat if c["feat"] else None), requires=requires)
        assert torch.equal(out["radii"], buf["radii"])
        fx = {}
        for k, v in scene.items():
            fx["in_" + k] = v.numpy()
        fx.update(in_viewmatrix=cam.viewmatrix.numpy(), in_projmatrix=cam.projmatrix.numpy(),
                  in_campos=cam.campos.numpy(), in_tanfov=np.array([cam.tanfovx, cam.tanfovy], np.float64),
                  in_size=np.array([c["W"], c["H"]], np.int32), in_bg=np.array(c["bg"], np.float32),
                  in_sh_degree=np.array(c["deg"], np.int32), in_include_feature=np.array(int(c["feat"]), np.int32),
                  in_feature_only=np.array(int(c["feature_only"]), np.int32), in_dL_dcolor=dcol.numpy(),
                  in_dL_dfeature=dfeat.numpy())
        fx["num_rendered"] = np.array(int(buf["num_rendered"]), np.int64)
        for k in ("color", "feature", "radii"):
            fx[k] = buf[k].detach().cpu().numpy()
        for k in ("depths", "means2D", "cov3D", "conic_opacity", "rgb", "tiles_touched", "point_list", "keys",
                  "final_T", "n_contrib", "ranges"):
            fx[k] = np.ascontiguousarray(buf[k]).copy()
        for k, g in out["grads"].items():
            if g is not None:
                fx["grad_" + k] = g.detach().cpu().numpy()
        path = os.path.join(out_dir, name + ".npz")
        np.savez_compressed(path, **fx)
        print(f"{name}: R={int(buf['num_rendered'])} V={int((fx['radii'] > 0).sum())} -> {path} "
              f"({os.path.getsize(path) / 1024:.0f} KiB)")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE))
