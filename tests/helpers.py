"""Shared helpers for the parity tests: run the same seeded scene through the CUDA path (via the public module,
hence through the C ABI), the CPU oracle, and — on the GPU box — the unmodified reference extension in baseline/_ref."""
from __future__ import annotations

import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from langsplat_b200 import synthetic as syn  # noqa: E402
from langsplat_b200 import rasterizer as R  # noqa: E402
from oracle import oracle as orc  # noqa: E402


def settings_for(cam, device, include_feature=True, sh_degree=3, bg=(0.0, 0.0, 0.0), scale_modifier=1.0, debug=False,
                 cls=R.GaussianRasterizationSettings):
    return cls(image_height=cam.image_height, image_width=cam.image_width, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
               bg=torch.tensor(bg, dtype=torch.float32, device=device), scale_modifier=scale_modifier,
               viewmatrix=cam.viewmatrix.to(device), projmatrix=cam.projmatrix.to(device), sh_degree=sh_degree,
               campos=cam.campos.to(device), prefiltered=False, debug=debug, include_feature=include_feature)


def oracle_forward(scene, cam, include_feature=True, sh_degree=3, bg=(0.0, 0.0, 0.0), scale_modifier=1.0,
                   colors_precomp=None, cov3D_precomp=None, dtype=np.float32):
    return orc.forward(
        scene["means3D"], scene["opacities"], cam.viewmatrix, cam.projmatrix, cam.campos, cam.image_width,
        cam.image_height, cam.tanfovx, cam.tanfovy, np.asarray(bg), shs=None if colors_precomp is not None else scene["shs"],
        colors_precomp=colors_precomp, language_feature=scene["language_feature"] if include_feature else None,
        scales=None if cov3D_precomp is not None else scene["scales"],
        rotations=None if cov3D_precomp is not None else scene["rotations"], cov3D_precomp=cov3D_precomp,
        scale_modifier=scale_modifier, sh_degree=sh_degree, include_feature=include_feature, dtype=dtype)


def cuda_forward_state(scene, cam, device="cuda", include_feature=True, sh_degree=3, bg=(0.0, 0.0, 0.0),
                       scale_modifier=1.0, colors_precomp=None, cov3D_precomp=None):
    """Forward through the product's internal entry (keeps the scratch so sorted keys can be inspected)."""
    rs = settings_for(cam, device, include_feature, sh_degree, bg, scale_modifier)
    d = {k: v.to(device) for k, v in scene.items()}
    color, feat, radii, state = R._forward_impl(
        rs, d["means3D"], None if colors_precomp is not None else d["shs"],
        None if colors_precomp is None else colors_precomp.to(device), d["language_feature"], d["opacities"],
        None if cov3D_precomp is not None else d["scales"], None if cov3D_precomp is not None else d["rotations"],
        None if cov3D_precomp is None else cov3D_precomp.to(device), keep_scratch=True)
    torch.cuda.synchronize()
    return color, feat, radii, state


def run_module(module_GR, module_GS, scene, cam, device="cuda", include_feature=True, sh_degree=3, bg=(0.0, 0.0, 0.0),
               grads=None, requires=("means3D", "shs", "language_feature", "opacities", "scales", "rotations"),
               scale_modifier=1.0, use_colors_precomp=None, use_cov3D_precomp=None):
    """Forward (+ backward if grads=(dcol, dfeat)) through a GaussianRasterizer-shaped public API."""
    rs = settings_for(cam, device, include_feature, sh_degree, bg, scale_modifier, cls=module_GS)
    t = {k: v.to(device).clone().requires_grad_(k in requires) for k, v in scene.items()}
    means2D = torch.zeros_like(t["means3D"], requires_grad=True)
    colors_precomp = cov3D_precomp = None
    if use_colors_precomp is not None:
        colors_precomp = use_colors_precomp.to(device).clone().requires_grad_("colors_precomp" in requires)
    if use_cov3D_precomp is not None:
        cov3D_precomp = use_cov3D_precomp.to(device).clone().requires_grad_("cov3D_precomp" in requires)
    rast = module_GR(raster_settings=rs)
    lang = t["language_feature"] if include_feature else torch.zeros((1,), dtype=torch.float32, device=device)
    color, feat, radii = rast(means3D=t["means3D"], means2D=means2D, shs=None if colors_precomp is not None else t["shs"],
                              colors_precomp=colors_precomp, language_feature_precomp=lang, opacities=t["opacities"],
                              scales=None if cov3D_precomp is not None else t["scales"],
                              rotations=None if cov3D_precomp is not None else t["rotations"],
                              cov3D_precomp=cov3D_precomp)
    out = dict(color=color.detach(), feature=feat.detach(), radii=radii.detach())
    if grads is not None:
        dcol, dfeat = grads
        loss = 0
        if dcol is not None:
            loss = loss + (color * dcol.to(device)).sum()
        if dfeat is not None and include_feature:
            loss = loss + (feat * dfeat.to(device)).sum()
        loss.backward()
        g = {k: (v.grad.detach() if v.grad is not None else None) for k, v in t.items()}
        g["means2D"] = means2D.grad.detach() if means2D.grad is not None else None
        if colors_precomp is not None:
            g["colors_precomp"] = colors_precomp.grad
        if cov3D_precomp is not None:
            g["cov3D_precomp"] = cov3D_precomp.grad
        out["grads"] = g
    torch.cuda.synchronize()
    return out


def load_reference():
    """The unmodified reference extension built into baseline/_ref; None if absent."""
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref_dir, "diff_gaussian_rasterization")):
        return None
    saved = sys.modules.pop("diff_gaussian_rasterization", None)
    saved_c = sys.modules.pop("diff_gaussian_rasterization._C", None)
    sys.path.insert(0, ref_dir)
    try:
        mod = importlib.import_module("diff_gaussian_rasterization")
        assert os.path.realpath(mod.__file__).startswith(os.path.realpath(ref_dir)), mod.__file__
    except Exception:
        mod = None
    finally:
        sys.path.remove(ref_dir)
        sys.modules.pop("diff_gaussian_rasterization", None)
        sys.modules.pop("diff_gaussian_rasterization._C", None)
        if saved is not None:
            sys.modules["diff_gaussian_rasterization"] = saved
        if saved_c is not None:
            sys.modules["diff_gaussian_rasterization._C"] = saved_c
    return mod


def _al(x, a=128):
    return (x + a - 1) // a * a


def ref_forward_buffers(REF, scene, cam, include_feature=True, sh_degree=3, bg=(0.0, 0.0, 0.0)):
    """Calls the reference extension's _C.rasterize_gaussians and parses its three opaque byte buffers
    (RAST/cuda_rasterizer/rasterizer_impl.cu:155-194, 128-byte `obtain` alignment)."""
    d = {k: v.cuda() for k, v in scene.items()}
    rs = settings_for(cam, "cuda", include_feature, sh_degree, bg)
    lang = d["language_feature"] if include_feature else torch.zeros((1,), device="cuda")
    e = torch.Tensor([])
    args = (rs.bg, d["means3D"], e, lang, d["opacities"], d["scales"], d["rotations"], rs.scale_modifier, e,
            rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, d["shs"],
            rs.sh_degree, rs.campos, rs.prefiltered, rs.debug, rs.include_feature)
    num_rendered, color, feat, radii, geomB, binB, imgB = REF._C.rasterize_gaussians(*args)
    torch.cuda.synchronize()
    P = d["means3D"].shape[0]
    N = cam.image_width * cam.image_height
    T = ((cam.image_width + 15) // 16) * ((cam.image_height + 15) // 16)
    geom = geomB.cpu().numpy()
    off = 0
    out = dict(num_rendered=num_rendered, color=color, feature=feat, radii=radii)

    def take(buf, off, dtype, count):
        off = _al(off)
        nbytes = np.dtype(dtype).itemsize * count
        return buf[off:off + nbytes].view(dtype), off + nbytes

    out["depths"], off = take(geom, off, np.float32, P)
    _, off = take(geom, off, np.uint8, 3 * P)
    _, off = take(geom, off, np.int32, P)
    out["means2D"], off = take(geom, off, np.float32, 2 * P)
    out["cov3D"], off = take(geom, off, np.float32, 6 * P)
    out["conic_opacity"], off = take(geom, off, np.float32, 4 * P)
    out["rgb"], off = take(geom, off, np.float32, 3 * P)
    out["tiles_touched"], off = take(geom, off, np.uint32, P)
    Rn = num_rendered
    binb = binB.cpu().numpy()
    off = 0
    out["point_list"], off = take(binb, off, np.uint32, Rn)
    _, off = take(binb, off, np.uint32, Rn)
    out["keys"], off = take(binb, off, np.uint64, Rn)
    img = imgB.cpu().numpy()
    off = 0
    out["final_T"], off = take(img, off, np.float32, N)
    out["n_contrib"], off = take(img, off, np.uint32, N)
    ranges, off = take(img, off, np.uint32, 2 * N)
    out["ranges"] = ranges[:2 * T].reshape(T, 2)
    return out


def _f64_pair(a, b):
    a = a.detach() if hasattr(a, "detach") else torch.as_tensor(a)
    b = b.detach() if hasattr(b, "detach") else torch.as_tensor(b)
    dev = a.device if a.is_cuda else b.device  # stay on the GPU when either side lives there (full-size tensors)
    return a.to(dev).double(), b.to(dev).double()


def rel_err(a, b):
    """max |a-b| / max |b| — the relative metric used for gradients (the reference sums with unordered atomics)."""
    a, b = _f64_pair(a, b)
    denom = b.abs().max().item() if b.numel() else 0.0
    return ((a - b).abs().max().item() if b.numel() else 0.0) / (denom if denom > 0 else 1.0)


def rel_l2(a, b):
    """||a-b||_2 / ||b||_2 — a second gradient metric that, unlike rel_err, is not set by the largest entry alone."""
    a, b = _f64_pair(a, b)
    denom = b.norm().item()
    return (a - b).norm().item() / (denom if denom > 0 else 1.0)


def record_errors(case: str, table: dict) -> None:
    """Appends the per-tensor errors a parity test measured to gpurun_out/grad_errors.json (copied into
    profiles/rNN_grad_errors.json after a GPU call, so the numbers behind the assertions are tracked)."""
    import json
    out_dir = os.path.join(ROOT, "gpurun_out")
    try:
        os.makedirs(out_dir, exist_ok=True)
        path = os.path.join(out_dir, "grad_errors.json")
        doc = json.load(open(path)) if os.path.exists(path) else {}
        doc[case] = table
        with open(path, "w") as f:
            json.dump(doc, f, indent=1, sort_keys=True)
    except OSError:
        pass


def compare_with_reference_runs(case: str, ours: dict, refs: list, keys, bar: float = 1e-4, arbiter=None) -> None:
    """ours / refs[i]: {name: gradient tensor}.  The reference sums with unordered float atomics, so one reference run
    is not a fixed point: compare with the mean of the runs and let 3x the runs' own scatter about that mean widen the
    bar where it is the larger quantity.  Both metrics (max-abs ratio, relative L2) of ours-vs-mean and
    of every reference run vs the mean go to the tracked error log.

    `arbiter` (optional): callable returning {name: float64 gradient} of the same step from the CPU oracle evaluated
    in float64.  Some Gaussians are ill-conditioned in float32 (needles a fraction of a pixel thin: config 4's Gaussian
    1772573 has dL_drotation = 11.43 in float64, 11.26 from the reference and 11.66 from this repo — both 2 % off, in
    opposite directions, while each implementation repeats itself to 0.5 %), so run-to-run scatter does not bound the
    distance between two correct float32 evaluations there.  Where an element misses the bar, float64 decides: this
    repo's value must be no farther from it than FOUR times the reference's distance (or the bar) — both are a few
    per cent off on such an element, the factor only separates that from a wrong result —, such elements must be
    rare (< 1e-5 of the tensor), and the relative-L2 bar must hold on the rest.  Every arbitrated element is logged
    with the three values (profiles/rNN_grad_errors.json)."""
    table, failures, truth = {}, [], None
    for k in keys:
        mean = sum(r[k].double() for r in refs) / len(refs)
        noise = max(rel_err(r[k], mean) for r in refs)
        noise_l2 = max(rel_l2(r[k], mean) for r in refs)
        err, err_l2 = rel_err(ours[k], mean), rel_l2(ours[k], mean)
        bar_max, bar_l2 = max(bar, 3.0 * noise), max(bar, 3.0 * noise_l2)
        table[k] = dict(ours_vs_mean_maxabs=err, ours_vs_mean_l2=err_l2, ref_vs_mean_maxabs=noise,
                        ref_vs_mean_l2=noise_l2, bar_maxabs=bar_max, bar_l2=bar_l2,
                        reference_runs=len(refs), max_abs_value=float(mean.abs().max()))
        if err < bar_max and err_l2 < bar_l2:
            continue
        if arbiter is None:
            if not err < bar_max:
                failures.append(f"{k}: max-abs rel err {err:.3e} (reference runs about their mean: {noise:.3e})")
            if not err_l2 < bar_l2:
                failures.append(f"{k}: rel L2 err {err_l2:.3e} (reference runs about their mean: {noise_l2:.3e})")
            continue
        if truth is None:
            truth = arbiter()
        t = torch.as_tensor(truth[k]).to(mean.device).double().reshape(mean.shape)
        o = ours[k].double()
        scale = mean.abs().max().item()
        out = (o - mean).abs() > bar_max * scale  # the elements that miss the bar
        n_out = int(out.sum())
        d_ours, d_ref = (o - t).abs()[out], (mean - t).abs()[out]
        worst = float((d_ours / torch.clamp(torch.maximum(4.0 * d_ref, torch.full_like(d_ref, bar * scale)),
                                            min=1e-300)).max()) if n_out else 0.0
        if n_out:  # the worst of them, for the log: flat index, this repo, reference mean, float64
            ratio = d_ours / torch.clamp(torch.maximum(4.0 * d_ref, torch.full_like(d_ref, bar * scale)), min=1e-300)
            top = ratio.topk(min(5, n_out))
            flat_idx = out.flatten().nonzero().flatten()[top.indices]
            table[k]["worst_elements"] = [dict(index=int(i), ours=float(o.flatten()[i]), reference=float(mean.flatten()[i]),
                                               float64=float(t.flatten()[i]), ratio=float(r))
                                          for i, r in zip(flat_idx.tolist(), top.values.tolist())]
        rest = ~out
        l2_rest = ((o - mean)[rest].norm() / mean[rest].norm()).item()
        table[k].update(arbitrated_elements=n_out, arbitrated_fraction=n_out / out.numel(),
                        worst_ours_over_4x_reference_distance_to_f64=worst, l2_without_arbitrated=l2_rest,
                        ours_vs_f64_maxabs=rel_err(o, t), ref_mean_vs_f64_maxabs=rel_err(mean, t))
        if n_out > max(8, 1e-5 * out.numel()):
            failures.append(f"{k}: {n_out} elements miss the bar {bar_max:.2e} (too many to be conditioning outliers)")
        if worst > 1.0:
            failures.append(f"{k}: an element is {worst:.2f}x farther from the float64 value than allowed "
                            f"(4x the reference's distance)")
        if not l2_rest < bar_l2:
            failures.append(f"{k}: rel L2 err {l2_rest:.3e} without the arbitrated elements")
    record_errors(case, table)
    assert not failures, "; ".join(failures)


def oracle_step_f64(scene, cams, grad_seed0):
    """The summed gradients of a multi-view step from the CPU oracle in float64 (upstream gradients
    syn.make_grads(.., grad_seed0 + view)); keys as run_module's `grads`."""
    total = {}
    names = dict(means3D="means3D", means2D="means2D", shs="shs", language_feature="language_feature",
                 opacities="opacities", scales="scales", rotations="rotations")
    for i, cam in enumerate(cams):
        dcol, dfeat = syn.make_grads(cam.image_width, cam.image_height, grad_seed0 + i)
        fw = oracle_forward(scene, cam, dtype=np.float64)
        g = orc.backward(fw, dcol.numpy(), dfeat.numpy())
        for k, ok in names.items():
            v = np.asarray(g[ok], np.float64)
            total[k] = v if k not in total else total[k] + v
    return total


# ---- golden vectors (tests/golden/*.npz, produced by tests/golden/make_golden.py from the reference extension) ----
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
GOLDEN_CASES = ("rgb_feat_bg", "feature_only_ragged", "rgb_only_deg1")


def load_golden(name):
    """Returns (fixture dict, scene dict of CPU tensors, Camera, kwargs for oracle_forward / run_module)."""
    path = os.path.join(GOLDEN_DIR, name + ".npz")
    fx = dict(np.load(path))
    scene = {k[3:]: torch.from_numpy(fx[k]) for k in ("in_means3D", "in_scales", "in_rotations", "in_opacities",
                                                      "in_shs", "in_language_feature")}
    W, H = int(fx["in_size"][0]), int(fx["in_size"][1])
    cam = syn.Camera(W, H, float(fx["in_tanfov"][0]), float(fx["in_tanfov"][1]), torch.from_numpy(fx["in_viewmatrix"]),
                     torch.from_numpy(fx["in_projmatrix"]), torch.from_numpy(fx["in_campos"]))
    kw = dict(include_feature=bool(fx["in_include_feature"]), sh_degree=int(fx["in_sh_degree"]),
              bg=tuple(float(x) for x in fx["in_bg"]))
    return fx, scene, cam, kw


def golden_available():
    return all(os.path.exists(os.path.join(GOLDEN_DIR, n + ".npz")) for n in GOLDEN_CASES)
