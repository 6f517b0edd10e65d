"""Times the fused decode + relevancy kernels on one 1440x1080 feature map (random-init weights of the reference's
decoder architecture): python profiles/tools/decoder_perf.py [precisions...]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from langsplat_b200.decode import FeatureDecoder  # noqa: E402

dev = torch.device("cuda:0")
dims = [3, 16, 32, 64, 128, 256, 256, 512]
g = torch.Generator().manual_seed(9)
ws = [(torch.randn(dims[i + 1], dims[i], generator=g) / dims[i] ** 0.5).to(dev) for i in range(7)]
bs = [(0.1 * torch.randn(dims[i + 1], generator=g)).to(dev) for i in range(7)]
ph = torch.randn(7, 512, generator=g)
ph = (ph / ph.norm(dim=-1, keepdim=True)).to(dev)
n = 1440 * 1080
x = torch.randn(n, 3, generator=g)
x = (x / x.norm(dim=-1, keepdim=True)).to(dev)
dec = FeatureDecoder(ws, bs)
dec.set_phrases(ph[:3], ph[3:])
ref = dec.relevancy_maps(x, precision="fp32")
for prec in (sys.argv[1:] or ["tf32", "f16"]):
    for what, fn in (("relevancy", lambda: dec.relevancy_maps(x, precision=prec)), ("decode", lambda: dec.decode(x, precision=prec))):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        print(f"{prec} {what}: {e0.elapsed_time(e1) / 10:.3f} ms", flush=True)
    print(prec, "max diff vs fp32 variant", float((dec.relevancy_maps(x, precision=prec) - ref).abs().max()), flush=True)

# per-phase clock stamps of CTA 0's second tile (cycles)
import ctypes  # noqa: E402
from langsplat_b200 import _lib  # noqa: E402
buf = torch.zeros(64, dtype=torch.int64, device=dev)
_lib.load().lsr_debug_decode_trace(ctypes.c_void_p(buf.data_ptr()))
for prec in (sys.argv[1:] or ["tf32", "f16"]):
    buf.zero_()
    dec.relevancy_maps(x, precision=prec)
    torch.cuda.synchronize()
    t = buf.cpu().tolist()
    k = max(i for i, v in enumerate(t) if v) + 1
    rows = [(t[1 + 3 * p] - (t[3 * p] if p == 0 else t[3 * p]), t[2 + 3 * p] - t[1 + 3 * p], t[3 + 3 * p] - t[2 + 3 * p]) for p in range((k - 1) // 3)]
    print(prec, "tile cycles", t[k - 1] - t[0], "per pass (issue, wait-for-accumulator, epilogue):", rows, flush=True)
_lib.load().lsr_debug_decode_trace(None)
