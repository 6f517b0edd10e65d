#!/usr/bin/env python
"""Where does the gradient error against the reference extension come from?  Runs config 4 (3 M Gaussians, a few
views), this repo's backward against the mean of several reference runs, per tensor (max-abs ratio and relative L2),
including the internal dL_dcov3D / dL_dmeans2D that the reference's _C.rasterize_gaussians_backward returns, and lists
the worst elements of dL_drotations with the attributes of their Gaussians.
Usage: python profiles/tools/grad_diag.py [--views 2] [--config 4] [--lib path/to/liblsr.so]"""
import argparse
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
ap = argparse.ArgumentParser()
ap.add_argument("--views", type=int, default=2)
ap.add_argument("--config", type=int, default=4)
ap.add_argument("--lib", default=None)
ap.add_argument("--ref-runs", type=int, default=3)
args = ap.parse_args()
if args.lib:
    shutil.copy(args.lib, os.path.join(ROOT, "langsplat_b200", "liblsr.so"))

import torch  # noqa: E402
from helpers import R, load_reference, rel_err, rel_l2, settings_for, syn  # noqa: E402

REF = load_reference()
dev = torch.device("cuda:0")
scene, cams = syn.make_config(args.config, n_cameras=args.views)
names = ("means3D", "shs", "opacities", "scales", "rotations", "language_feature")


def step(GR, GS):
    params = {k: scene[k].to(dev).requires_grad_(True) for k in names}
    means2D = torch.zeros_like(params["means3D"], requires_grad=True)
    for i, cam in enumerate(cams):
        dcol, dfeat = [t.to(dev) for t in syn.make_grads(cam.image_width, cam.image_height, 2400 + i)]
        rast = GR(raster_settings=settings_for(cam, dev, cls=GS))
        color, feat, radii = rast(means3D=params["means3D"], means2D=means2D, shs=params["shs"],
                                  language_feature_precomp=params["language_feature"], opacities=params["opacities"],
                                  scales=params["scales"], rotations=params["rotations"])
        torch.autograd.backward([color, feat], [dcol, dfeat])
    torch.cuda.synchronize()
    out = {k: p.grad.detach() for k, p in params.items()}
    out["means2D"] = means2D.grad.detach()
    return out, radii


refs = [step(REF.GaussianRasterizer, REF.GaussianRasterizationSettings)[0] for _ in range(args.ref_runs)]
ours = [step(R.GaussianRasterizer, R.GaussianRasterizationSettings) for _ in range(2)]
radii = ours[0][1]
ours = [o[0] for o in ours]
print(f"lib={args.lib or 'in-tree'} config={args.config} views={args.views}")
for k in list(names) + ["means2D"]:
    mean = sum(r[k].double() for r in refs) / len(refs)
    print(f"  {k:18s} ours-vs-mean max {rel_err(ours[0][k], mean):.2e} l2 {rel_l2(ours[0][k], mean):.2e} | "
          f"ours-vs-ours max {rel_err(ours[0][k], ours[1][k]):.2e} | ref-vs-mean max "
          f"{max(rel_err(r[k], mean) for r in refs):.2e} l2 {max(rel_l2(r[k], mean) for r in refs):.2e}")
k = "rotations"
mean = sum(r[k].double() for r in refs) / len(refs)
err = (ours[0][k].double() - mean).abs()
scale = mean.abs().max()
flat = err.flatten().topk(8)
sc, op = scene["scales"], scene["opacities"]
for v, idx in zip(flat.values.tolist(), flat.indices.tolist()):
    g, c = idx // 4, idx % 4
    print(f"    gaussian {g} comp {c}: err/max {v / scale:.2e} ours {ours[0][k][g, c].item():.6g} {ours[1][k][g, c].item():.6g} "
          f"ref {[round(r[k][g, c].item(), 6) for r in refs]} scales {sc[g].tolist()} opacity {op[g].item():.3f} "
          f"radius(last view) {radii[g].item()}")
