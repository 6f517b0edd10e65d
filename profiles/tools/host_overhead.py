"""Host-side overhead of one forward+backward through the public module (tiny scene, so GPU time is negligible)."""
import sys, os, time, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from langsplat_b200 import GaussianRasterizationSettings, GaussianRasterizer, synthetic as syn
dev = torch.device("cuda:0")
scene, cams = syn.make_config(1, P=2000, W=64, H=64)
cam = cams[0]
rs = GaussianRasterizationSettings(cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, torch.zeros(3, device=dev),
                                   1.0, cam.viewmatrix.to(dev), cam.projmatrix.to(dev), 3, cam.campos.to(dev), False, False, True)
t = {k: v.to(dev).requires_grad_(True) for k, v in scene.items()}
m2 = torch.zeros_like(t["means3D"], requires_grad=True)
dc, df = [x.to(dev) for x in syn.make_grads(64, 64, 1)]
rast = GaussianRasterizer(rs)
def step(bwd=True):
    c, f, r = rast(means3D=t["means3D"], means2D=m2, shs=t["shs"], language_feature_precomp=t["language_feature"],
                   opacities=t["opacities"], scales=t["scales"], rotations=t["rotations"])
    if bwd:
        torch.autograd.backward([c, f], [dc, df])
for _ in range(20): step()
torch.cuda.synchronize()
for bwd in (False, True):
    n = 300
    t0 = time.perf_counter()
    for _ in range(n): step(bwd)
    torch.cuda.synchronize()
    print(("fwd+bwd" if bwd else "fwd only"), "%.1f us per call (host-bound, tiny scene)" % ((time.perf_counter() - t0) / n * 1e6))
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(200): step()
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
