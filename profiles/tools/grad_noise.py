"""Gradient agreement at full size (config 2): ours vs the reference extension, and each against a second run of
itself (both sum with unordered float atomics).  rel err = max-abs difference / max-abs value, per tensor."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from helpers import load_reference, rel_err, run_module
from langsplat_b200 import rasterizer as R, synthetic as syn
REF = load_reference()
scene, cams = syn.make_config(int(sys.argv[1]) if len(sys.argv) > 1 else 2)
cam = cams[0]
dcol, dfeat = syn.make_grads(cam.image_width, cam.image_height, 2002)
run = lambda m: run_module(m.GaussianRasterizer, m.GaussianRasterizationSettings, scene, cam, grads=(dcol, dfeat))["grads"]
o1, o2 = run(R), run(R)
r1, r2 = (run(REF), run(REF)) if REF is not None else (None, None)
truth = None
if os.environ.get("F64", "1") == "1":
    import numpy as np
    from oracle import oracle as orc
    fw = orc.forward(scene["means3D"], scene["opacities"], cam.viewmatrix, cam.projmatrix, cam.campos, cam.image_width,
                     cam.image_height, cam.tanfovx, cam.tanfovy, np.zeros(3), shs=scene["shs"],
                     language_feature=scene["language_feature"], scales=scene["scales"], rotations=scene["rotations"],
                     dtype=np.float64)
    g = orc.backward(fw, dcol, dfeat)
    truth = {k: torch.from_numpy(np.asarray(g[k], np.float64)) for k in g}
for k in ("means3D", "means2D", "shs", "language_feature", "opacities", "scales", "rotations"):
    line = f"{k:18s} ours-ours {rel_err(o1[k], o2[k]):.2e}"
    if r1 is not None:
        line += f"  ref-ref {rel_err(r1[k], r2[k]):.2e}  ours-ref {rel_err(o1[k], r1[k]):.2e} {rel_err(o2[k], r2[k]):.2e}"
        d = (o1[k] - r1[k]).abs().flatten(); i = int(d.argmax())
        line += f"  worst@{i}: ours {o1[k].flatten()[i].item():.6e} ref {r1[k].flatten()[i].item():.6e} max|ref| {r1[k].abs().max().item():.3e}"
    if truth is not None and k in truth:
        t = truth[k].reshape(o1[k].shape)
        e = lambda a: ((a.double().cpu() - t).abs().max() / t.abs().max()).item()
        line += f"  | vs f64 oracle: ours {e(o1[k]):.2e}" + (f" ref {e(r1[k]):.2e}" if r1 is not None else "")
    print(line)
