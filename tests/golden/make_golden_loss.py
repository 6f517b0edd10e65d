"""Golden vectors for the fused L1 loss: inputs + outputs of the REFERENCE's own `l1_loss`
(/root/reference/utils/loss_utils.py:17-18) called the way train.py:96-102 calls it, evaluated with PyTorch on the CPU
in this container, gradients by autograd.  Run here (the reference is not present on the GPU box):
    python tests/golden/make_golden_loss.py
"""
import importlib.util
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("ref_loss_utils", "/root/reference/utils/loss_utils.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

out = {}
cases = [("rgb_33x17", (3, 17, 33), None, 1.0), ("feat_masked_40x24", (3, 24, 40), "hw", 1.0),
         ("feat_masked_odd_19x7", (3, 7, 19), "hw", 0.37), ("ties_16x16", (3, 16, 16), "hw", 2.0)]
for name, shape, mask_kind, gscale in cases:
    g = torch.Generator().manual_seed(sum(map(ord, name)))
    x = torch.rand(shape, generator=g, requires_grad=True)
    gt = torch.rand(shape, generator=g)
    mask = None
    if mask_kind:
        mask = (torch.rand((1,) + shape[1:], generator=g) > 0.3).float()
    if name.startswith("ties"):  # exact ties exercise sign(0) = 0
        with torch.no_grad():
            x[:, ::2, ::3] = gt[:, ::2, ::3]
    loss = ref.l1_loss(x, gt) if mask is None else ref.l1_loss(x * mask, gt * mask)
    (loss * gscale).backward()
    out[name + ".x"] = x.detach().numpy()
    out[name + ".gt"] = gt.numpy()
    if mask is not None:
        out[name + ".mask"] = mask.numpy()
    out[name + ".gscale"] = np.float32(gscale)
    out[name + ".loss"] = loss.detach().numpy()
    out[name + ".grad"] = x.grad.numpy()
np.savez_compressed(os.path.join(HERE, "l1_loss.npz"), **out)
print("wrote", os.path.join(HERE, "l1_loss.npz"), {k: v.shape for k, v in out.items() if k.endswith(".x")})
