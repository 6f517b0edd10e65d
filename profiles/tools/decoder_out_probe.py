import os, sys, torch
sys.path.insert(0, os.getcwd())
from langsplat_b200.decode import FeatureDecoder
dev = torch.device("cuda:0")
dims = [3, 16, 32, 64, 128, 256, 256, 512]
g = torch.Generator().manual_seed(9)
ws = [(torch.randn(dims[i + 1], dims[i], generator=g) / dims[i] ** 0.5).to(dev) for i in range(7)]
bs = [(0.1 * torch.randn(dims[i + 1], generator=g)).to(dev) for i in range(7)]
x = torch.randn(1440 * 1080, 3, generator=g).to(dev)
dec = FeatureDecoder(ws, bs, "f16")
for _ in range(3):
    y = dec.decode(x)
torch.cuda.synchronize()
