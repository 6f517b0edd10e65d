"""Golden vectors for the ground-truth language-feature map: inputs + outputs of the REFERENCE's own
Camera.get_language_feature (/root/reference/scene/cameras.py:58-92), called unbound on a stand-in object that carries
the three attributes the method reads (image_name, image_height, image_width) — Camera.__init__ itself needs a GPU.
The method ends in `.cuda()`; this container has no GPU, so Tensor.cuda is made the identity for the call.
Run here (the reference is not present on the GPU box):
    python tests/golden/make_golden_gt_features.py
"""
import os
import sys
import tempfile
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
import importlib.util  # noqa: E402

sys.path.insert(0, "/root/reference")  # cameras.py imports utils.graphics_utils; the file itself is loaded directly so
spec = importlib.util.spec_from_file_location("ref_cameras", "/root/reference/scene/cameras.py")  # that the package's
ref_cameras = importlib.util.module_from_spec(spec)  # __init__ (dataset readers, plyfile) is not pulled in
spec.loader.exec_module(ref_cameras)
Camera = ref_cameras.Camera

torch.Tensor.cuda = lambda self, *a, **k: self

out = {}
with tempfile.TemporaryDirectory() as d:
    for name, (Hs, Ws), (H, W), K, seg_dtype in (("full_37x53", (37, 53), (37, 53), 19, np.float32),
                                                  ("crop_40x64", (48, 70), (40, 64), 7, np.float32),
                                                  ("int_24x31", (24, 31), (24, 31), 5, np.int32)):
        rng = np.random.default_rng(sum(map(ord, name)))
        seg = rng.integers(-1, K, size=(4, Hs, Ws)).astype(seg_dtype)
        seg[0, :3] = -1          # a band without any mask
        seg[2] = -1              # a level without any mask at all
        feat = rng.standard_normal((K, 3)).astype(np.float32)
        np.save(os.path.join(d, name + "_s.npy"), seg)
        np.save(os.path.join(d, name + "_f.npy"), feat)
        cam = types.SimpleNamespace(image_name=name, image_height=H, image_width=W)
        out[f"{name}.seg"], out[f"{name}.table"] = seg, feat
        out[f"{name}.hw"] = np.array([H, W], np.int32)
        for level in range(4):
            pf, mask = Camera.get_language_feature(cam, d, level)
            out[f"{name}.l{level}.feature"] = pf.contiguous().numpy()
            out[f"{name}.l{level}.mask"] = mask.numpy()
np.savez_compressed(os.path.join(HERE, "gt_features.npz"), **out)
print("wrote", os.path.join(HERE, "gt_features.npz"), len(out), "arrays")
