"""Golden vectors for the fused activations: inputs + outputs of the REFERENCE's own GaussianModel properties
(/root/reference/scene/gaussian_model.py:28-41 setup_functions, :134-155 get_scaling / get_rotation / get_opacity) and
of the renderer's feature normalisation (gaussian_renderer/__init__.py:86-88, the expression is inline in render(),
restated here verbatim), evaluated with PyTorch on the CPU in this container; gradients by autograd.
The model file imports two packages that are absent here and unused by these properties (simple_knn, plyfile): they
are stubbed for the import only.  Run here (the reference is not present on the GPU box):
    python tests/golden/make_golden_activations.py
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
for name, attrs in (("simple_knn", {}), ("simple_knn._C", {"distCUDA2": None}),
                    ("plyfile", {"PlyData": None, "PlyElement": None})):
    if name not in sys.modules:
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
sys.path.insert(0, "/root/reference")
from scene.gaussian_model import GaussianModel  # noqa: E402

out = {}
for name, P, seed in (("mixed_257", 257, 1), ("edge_64", 64, 2)):
    g = torch.Generator().manual_seed(seed)
    pc = GaussianModel(3)
    scaling = (torch.randn(P, 3, generator=g) * 1.5 - 3.0)
    rotation = torch.randn(P, 4, generator=g)
    opacity = torch.randn(P, 1, generator=g) * 4.0
    feature = torch.randn(P, 3, generator=g)
    if name.startswith("edge"):
        rotation[0] = 0.0            # ||q|| < eps: clamp_min branch
        rotation[1] = 1e-13          # still below eps
        rotation[2] = torch.tensor([1e-7, 0.0, 0.0, 0.0])
        feature[0] = 0.0             # ||f|| = 0: only the 1e-9 remains in the denominator
        feature[1] = 1e-10
        opacity[0], opacity[1], opacity[2] = 40.0, -40.0, -100.0
        scaling[0], scaling[1] = 10.0, -60.0
    pc._scaling = scaling.clone().requires_grad_(True)
    pc._rotation = rotation.clone().requires_grad_(True)
    pc._opacity = opacity.clone().requires_grad_(True)
    pc._language_feature = feature.clone().requires_grad_(True)
    scales, rots, opac = pc.get_scaling, pc.get_rotation, pc.get_opacity
    language_feature_precomp = pc.get_language_feature
    language_feature_precomp = language_feature_precomp / (language_feature_precomp.norm(dim=-1, keepdim=True) + 1e-9)
    ups = [torch.randn(t.shape, generator=g) for t in (scales, rots, opac, language_feature_precomp)]
    torch.autograd.backward([scales, rots, opac, language_feature_precomp], ups)
    for key, raw, y, up in (("scaling", pc._scaling, scales, ups[0]), ("rotation", pc._rotation, rots, ups[1]),
                            ("opacity", pc._opacity, opac, ups[2]),
                            ("feature", pc._language_feature, language_feature_precomp, ups[3])):
        out[f"{name}.{key}.raw"] = raw.detach().numpy()
        out[f"{name}.{key}.out"] = y.detach().numpy()
        out[f"{name}.{key}.up"] = up.numpy()
        out[f"{name}.{key}.grad"] = raw.grad.numpy()
np.savez_compressed(os.path.join(HERE, "activations.npz"), **out)
print("wrote", os.path.join(HERE, "activations.npz"), len(out), "arrays")
