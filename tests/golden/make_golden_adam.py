"""Golden vectors for the fused Adam step: torch.optim.Adam configured the way the reference configures it
(/root/reference/scene/gaussian_model.py:218-227: six groups with their own lr, `Adam(l, lr=0.0, eps=1e-15)`), run on
the CPU in this container for three optimizer.step() calls with seeded gradients (GaussianModel.training_setup itself
allocates on "cuda" and cannot run here; the learning rates are arguments/__init__.py:74-84's defaults).
    python tests/golden/make_golden_adam.py
"""
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
P = 193
g = torch.Generator().manual_seed(42)
shapes = {"xyz": (P, 3), "f_dc": (P, 1, 3), "f_rest": (P, 15, 3), "opacity": (P, 1), "scaling": (P, 3), "rotation": (P, 4)}
lrs = {"xyz": 0.00016 * 4.7, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001}
params = {k: torch.nn.Parameter(torch.randn(s, generator=g)) for k, s in shapes.items()}
l = [{"params": [params[k]], "lr": lrs[k], "name": k} for k in shapes]
opt = torch.optim.Adam(l, lr=0.0, eps=1e-15)
out = {"names": np.array(list(shapes)), "lrs": np.array([lrs[k] for k in shapes], np.float64)}
for k in shapes:
    out[f"{k}.p0"] = params[k].detach().numpy().copy()
for step in range(1, 4):
    for k in shapes:
        grad = torch.randn(shapes[k], generator=g) * (10.0 ** float(torch.randint(-6, 1, (1,), generator=g)))
        if step == 2:
            grad[::3] = 0.0  # Gaussians that were not visible in that view: zero gradient, moments still decay
        params[k].grad = grad
        out[f"{k}.g{step}"] = grad.numpy().copy()
    opt.step()
    for k in shapes:
        st = opt.state[params[k]]
        out[f"{k}.p{step}"] = params[k].detach().numpy().copy()
        out[f"{k}.m{step}"] = st["exp_avg"].numpy().copy()
        out[f"{k}.v{step}"] = st["exp_avg_sq"].numpy().copy()
np.savez_compressed(os.path.join(HERE, "adam.npz"), **out)
print("wrote", os.path.join(HERE, "adam.npz"), len(out), "arrays")
