"""Max-abs difference of the rendered images (colour, language feature) between this repo's rasterizer and the
unmodified reference extension (baseline/_ref) at full size: python profiles/tools/image_diff.py [config ...]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import R, load_reference, run_module, syn  # noqa: E402

REF = load_reference()
for cfg in [int(a) for a in sys.argv[1:]] or [2, 3]:
    scene, cams = syn.make_config(cfg)
    ours = run_module(R.GaussianRasterizer, R.GaussianRasterizationSettings, scene, cams[0])
    ref = run_module(REF.GaussianRasterizer, REF.GaussianRasterizationSettings, scene, cams[0])
    dc = (ours["color"] - ref["color"]).abs()
    df = (ours["feature"] - ref["feature"]).abs()
    print("config %d: colour max-abs diff %.3e (mean %.3e, %d of %d pixels differ), feature %.3e (mean %.3e), radii equal %s"
          % (cfg, dc.max().item(), dc.mean().item(), int((dc > 0).sum()), dc.numel(), df.max().item(), df.mean().item(),
             bool(torch.equal(ours["radii"], ref["radii"]))), flush=True)
