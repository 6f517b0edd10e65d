"""Host-side timeline of the end-to-end step (C2 sizes): where the host spends its time between loss.item() and the
first kernel of the next forward, and how long each Python phase takes.  Prints averages over N steps."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from langsplat_b200 import GaussianRasterizationSettings as GS, GaussianRasterizer as GR, l1_loss, synthetic as syn

dev = torch.device("cuda:0")
scene, cams = syn.make_config(2)
cam = cams[0]
P, W, H = scene["means3D"].shape[0], cam.image_width, cam.image_height
rs = GS(H, W, cam.tanfovx, cam.tanfovy, torch.zeros(3, device=dev), 1.0, cam.viewmatrix.to(dev), cam.projmatrix.to(dev),
        3, cam.campos.to(dev), False, False, True)
names = ["means3D", "shs", "opacities", "scales", "rotations", "language_feature"]
params = {k: scene[k].to(dev).requires_grad_(True) for k in names}
means2D = torch.zeros((P, 3), device=dev, requires_grad=True)
gt_c, gt_f = torch.rand(3, H, W, device=dev), torch.rand(3, H, W, device=dev)
acc = {}
def mark(name, t0):
    t1 = time.perf_counter(); acc[name] = acc.get(name, 0.0) + (t1 - t0); return t1
def step():
    t = time.perf_counter()
    for p in params.values(): p.grad = None
    means2D.grad = None
    t = mark("zero_grads", t)
    rast = GR(raster_settings=rs)
    t = mark("construct", t)
    color, feat, radii = rast(means3D=params["means3D"], means2D=means2D, shs=params["shs"],
                              language_feature_precomp=params["language_feature"], opacities=params["opacities"],
                              scales=params["scales"], rotations=params["rotations"])
    t = mark("forward_call(incl. mid sync)", t)
    loss = l1_loss(color, gt_c) + l1_loss(feat, gt_f)
    t = mark("loss", t)
    loss.backward()
    t = mark("backward_call", t)
    v = loss.item()
    t = mark("item(wait for GPU)", t)
for _ in range(10): step()
acc.clear()
N = 50
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(N): step()
torch.cuda.synchronize(); total = (time.perf_counter() - t0) / N
print("step %.1f us" % (total * 1e6))
for k, v in acc.items(): print("  %-32s %.1f us" % (k, v / N * 1e6))
# inside the forward call: host time until lsr_forward is entered
import langsplat_b200.rasterizer as RZ
lib = RZ.L.load()
orig = lib.lsr_forward
tt = {"pre": 0.0, "c": 0.0, "n": 0}
class Wrap:
    def __call__(self, *a):
        t1 = time.perf_counter(); tt["pre"] += t1 - tt["t0"]
        r = orig(*a); tt["c"] += time.perf_counter() - t1; tt["n"] += 1; return r
lib.lsr_forward = Wrap()
for _ in range(N):
    for p in params.values(): p.grad = None
    tt["t0"] = time.perf_counter()
    color, feat, radii = GR(raster_settings=rs)(means3D=params["means3D"], means2D=means2D, shs=params["shs"],
                              language_feature_precomp=params["language_feature"], opacities=params["opacities"],
                              scales=params["scales"], rotations=params["rotations"])
    torch.cuda.synchronize()
print("forward call: python before lsr_forward %.1f us, inside lsr_forward (GPU idle at entry) %.1f us" % (tt["pre"] / N * 1e6, tt["c"] / N * 1e6))
