"""GPU-box diagnostic for the decode kernels (not a pytest file): runs the fp32 and the tcgen05 variants on the golden
vectors, prints the errors against the reference vectors and timings at an eval-image size.
    python tests/decoder_check.py
"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from langsplat_b200.decode import FeatureDecoder  # noqa: E402

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "decoder.npz"))
dev = torch.device("cuda:0")
L = int(G["n_layers"])
ws = [torch.from_numpy(G[f"w{i}"]).to(dev) for i in range(L)]
bs = [torch.from_numpy(G[f"b{i}"]).to(dev) for i in range(L)]
n_pos = int(G["n_pos"])
ph = torch.from_numpy(G["phrases"]).to(dev)
x = torch.from_numpy(G["x"]).to(dev)
# staged shapes: which layer shape breaks first if the tensor-core path is wrong
gen = torch.Generator().manual_seed(3)
for dims in ([3, 16], [8, 16], [16, 32], [3, 16, 32], [3, 256], [3, 256, 256], [3, 64, 512], [3, 16, 32, 64, 128]):
    w_ = [(torch.randn(dims[i + 1], dims[i], generator=gen) / dims[i] ** 0.5).to(dev) for i in range(len(dims) - 1)]
    b_ = [(0.1 * torch.randn(dims[i + 1], generator=gen)).to(dev) for i in range(len(dims) - 1)]
    xs = torch.randn(300, dims[0], generator=gen).to(dev)
    try:
        d_ = FeatureDecoder(w_, b_, "tf32")
        y32 = d_.decode(xs, precision="fp32")
        ytc = d_.decode(xs)
        yh = d_.decode(xs, precision="f16")
        torch.cuda.synchronize()
        print("staged", dims, "tf32 vs fp32 max err", float((ytc - y32).abs().max()), "f16 vs fp32",
              float((yh - y32).abs().max()), flush=True)
    except Exception as e:  # noqa: BLE001
        print("staged", dims, "FAILED", repr(e), flush=True)
        break
for prec in ("fp32", "tf32", "f16"):
    dec = FeatureDecoder(ws, bs, prec)
    dec.set_phrases(ph[:n_pos], ph[n_pos:])
    y = dec.decode(x)
    r = dec.relevancy_maps(x)
    torch.cuda.synchronize()
    print(prec, "decode err", float(np.abs(y.cpu().numpy() - G["y"]).max()),
          "relevancy err", float(np.abs(r.cpu().numpy() - G["relevancy"][..., 0]).max()), flush=True)
    big = torch.randn(1080 * 1440, 3, device=dev)
    big = big / big.norm(dim=-1, keepdim=True)
    for what, fn in (("relevancy", lambda: dec.relevancy_maps(big)), ("decode", lambda: dec.decode(big))):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            fn()
        e1.record()
        torch.cuda.synchronize()
        print(prec, what, "1440x1080:", e0.elapsed_time(e1) / 5, "ms", flush=True)
# the reference's way: nn.Linear chain + mm + softmax in PyTorch on the same GPU
with torch.no_grad():
    def ref(xx):
        h = xx
        for i in range(L):
            if i:
                h = torch.relu(h)
            h = torch.nn.functional.linear(h, ws[i], bs[i])
        h = h / h.norm(dim=-1, keepdim=True)
        out = torch.mm(h, ph.T)
        res = []
        for j in range(n_pos):
            pos = out[:, j:j + 1].repeat(1, ph.shape[0] - n_pos)
            sm = torch.softmax(10 * torch.stack((pos, out[:, n_pos:]), -1), -1)
            best = sm[..., 0].argmin(1)
            res.append(torch.gather(sm, 1, best[..., None, None].expand(best.shape[0], ph.shape[0] - n_pos, 2))[:, 0, 0])
        return torch.stack(res)
    big = torch.randn(1080 * 1440, 3, device=dev)
    ref(big)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        ref(big)
    torch.cuda.synchronize()
    print("torch (reference call sequence) relevancy 1440x1080:", (time.perf_counter() - t0) / 3 * 1e3, "ms", flush=True)
