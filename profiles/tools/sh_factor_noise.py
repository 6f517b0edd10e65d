"""Prints, per gradient, the run-to-run difference of two dense steps next to factored-vs-dense (3 views, degree d)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "tests"))
import torch
from helpers import rel_err, syn
import test_gpu_sh_factors as T

dev = torch.device("cuda:0")
scene, cams = syn.make_config(4, P=20000, W=200, H=136, n_cameras=3)
for d in (0, 1, 2, 3):
    a = T._step(scene, cams, dev, d, False)
    b = T._step(scene, cams, dev, d, False)
    f = T._step(scene, cams, dev, d, True)
    print("degree", d, " ".join("%s %.1e/%.1e" % (k[:6], rel_err(b[k], a[k]), rel_err(f[k], a[k])) for k in T.NAMES))
