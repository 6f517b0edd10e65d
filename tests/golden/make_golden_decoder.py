"""Golden vectors for the eval-side decode: the reference's own Autoencoder.decode
(/root/reference/autoencoder/model.py) with the dims eval/evaluate_iou_loc.py uses by default
(encoder 256,128,64,32,3 / decoder 16,32,64,128,256,256,512), random-init weights (seeded), run on the CPU in this
container on seeded 3-d features, and OpenCLIPNetwork.get_relevancy (/root/reference/eval/openclip_encoder.py:42-56)
called unbound on a stand-in object that carries seeded unit phrase embeddings (open_clip itself is not installed; the
method only touches pos_embeds / neg_embeds / positives / negatives).
    python tests/golden/make_golden_decoder.py
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference/autoencoder")
sys.path.insert(0, "/root/reference/eval")
for name in ("open_clip", "torchvision"):
    sys.modules.setdefault(name, types.ModuleType(name))
from model import Autoencoder  # noqa: E402
from openclip_encoder import OpenCLIPNetwork  # noqa: E402

torch.manual_seed(7)
model = Autoencoder([256, 128, 64, 32, 3], [16, 32, 64, 128, 256, 256, 512]).eval()
N, n_pos, n_neg = 301, 3, 4
g = torch.Generator().manual_seed(11)
x = torch.randn(N, 3, generator=g)
x = x / x.norm(dim=-1, keepdim=True)          # rendered features are (about) unit 3-vectors
x[5] *= 0.3
x[6] *= 1.7                                    # ... blended ones are not exactly
with torch.no_grad():
    y = model.decode(x)
clip = types.SimpleNamespace()
emb = torch.randn(n_pos + n_neg, 512, generator=g)
emb[0] = emb[0] * 0.5 + y[:50].mean(0) * 20   # one phrase that actually correlates with some pixels
emb = emb / emb.norm(dim=-1, keepdim=True)
clip.pos_embeds, clip.neg_embeds = emb[:n_pos], emb[n_pos:]
clip.positives, clip.negatives = ("a",) * n_pos, ("object", "things", "stuff", "texture")
with torch.no_grad():
    rel = torch.stack([OpenCLIPNetwork.get_relevancy(clip, y, j) for j in range(n_pos)])
out = {"x": x.numpy(), "y": y.numpy(), "phrases": emb.numpy(), "n_pos": np.int32(n_pos), "relevancy": rel.numpy()}
lin = [m for m in model.decoder if isinstance(m, torch.nn.Linear)]
for i, m in enumerate(lin):
    out[f"w{i}"] = m.weight.detach().numpy()
    out[f"b{i}"] = m.bias.detach().numpy()
out["n_layers"] = np.int32(len(lin))
np.savez_compressed(os.path.join(HERE, "decoder.npz"), **out)
print("wrote decoder.npz", y.shape, rel.shape, float(rel[..., 0].min()), float(rel[..., 0].max()))
