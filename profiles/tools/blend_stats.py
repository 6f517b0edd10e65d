"""Counts (footprint, instance) pairs for several warp footprints on a bench config (uses the CPU oracle's forward)."""
import ctypes, sys, os, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from langsplat_b200 import synthetic as syn
from oracle import oracle as orc
cfg = int(sys.argv[1]) if len(sys.argv) > 1 else 2
lib = ctypes.CDLL(sys.argv[2] if len(sys.argv) > 2 else "/tmp/libblendstats.so")
scene, cams = syn.make_config(cfg)
cam = cams[0]
fw = orc.forward(scene["means3D"], scene["opacities"], cam.viewmatrix, cam.projmatrix, cam.campos, cam.image_width,
                 cam.image_height, cam.tanfovx, cam.tanfovy, np.zeros(3), shs=scene["shs"],
                 language_feature=scene["language_feature"], scales=scene["scales"], rotations=scene["rotations"])
print("R", fw["num_rendered"], "K_exam", fw["k_exam"], "K_contrib", fw["k_contrib"])
p = lambda a: a.ctypes.data_as(ctypes.c_void_p)
for (w, h) in [(8, 4), (8, 8), (16, 8), (16, 16)]:
    out = np.zeros(10, np.int64)
    lib.blend_stats(cam.image_width, cam.image_height, p(fw["ranges"]), p(fw["point_list"]), p(fw["means2D"]),
                    p(fw["conic_opacity"]), p(fw["n_contrib"]), w, h, p(out))
    print(f"footprint {w}x{h}: fetched {out[0]} exam(tile-max) {out[1]} exam(fp-max) {out[2]} working pairs {out[3]} "
          f"lanes {out[4]} avg px/working pair {out[4] / max(out[3], 1):.1f} hist(1-2,3-4,5-8,9-16,17+) {out[5:10]}")
