"""ctypes front end of the CPU oracle (oracle/oracle.cpp).

TEST INFRASTRUCTURE ONLY — see the header of oracle.cpp.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module; the product package
(langsplat_b200) never does.

`forward()` / `backward()` mirror CudaRasterizer::Rasterizer::forward/backward
(RAST/cuda_rasterizer/rasterizer_impl.cu:198-355, 359-463) and return every intermediate the
reference keeps in its geometry / binning / image buffers so tests can compare stage by stage.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Dict, Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None


def build(force: bool = False) -> str:
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(
            os.path.join(_HERE, "oracle.cpp")):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.orc_preprocess_f32.restype = ctypes.c_int64
        _lib.orc_preprocess_f64.restype = ctypes.c_int64
        _lib.orc_higher_msb.restype = ctypes.c_uint32
        _lib.orc_num_threads.restype = ctypes.c_int
    return _lib


def num_threads() -> int:
    return int(lib().orc_num_threads())


def set_num_threads(n: int) -> None:
    lib().orc_set_num_threads(ctypes.c_int(n))


def higher_msb(n: int) -> int:
    return int(lib().orc_higher_msb(ctypes.c_uint32(n)))


def _np(x, dtype):
    if x is None:
        return None
    if hasattr(x, "detach"):
        x = x.detach().cpu().numpy()
    return np.ascontiguousarray(x, dtype=dtype)


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def forward(means3D, opacities, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy, bg, shs=None,
            colors_precomp=None, language_feature=None, scales=None, rotations=None, cov3D_precomp=None,
            scale_modifier=1.0, sh_degree=3, include_feature=True, dtype=np.float32) -> Dict[str, np.ndarray]:
    L = lib()
    suf = "f32" if dtype == np.float32 else "f64"
    real = ctypes.c_float if dtype == np.float32 else ctypes.c_double
    means3D = _np(means3D, dtype)
    P = means3D.shape[0]
    opacities = _np(opacities, dtype).reshape(-1)
    shs = _np(shs, dtype)
    colors_precomp = _np(colors_precomp, dtype)
    language_feature = _np(language_feature, dtype)
    scales = _np(scales, dtype)
    rotations = _np(rotations, dtype)
    cov3D_precomp = _np(cov3D_precomp, dtype)
    viewmatrix = _np(viewmatrix, dtype).reshape(-1)
    projmatrix = _np(projmatrix, dtype).reshape(-1)
    campos = _np(campos, dtype).reshape(-1)
    bg = _np(bg, dtype).reshape(-1)
    M = 0 if shs is None else shs.shape[1]
    # the C++ signatures narrow tanfov / scale_modifier to f32 (rasterize_points.cu:44,48-49)
    if dtype == np.float32:
        tanfovx, tanfovy, scale_modifier = float(np.float32(tanfovx)), float(np.float32(tanfovy)), float(
            np.float32(scale_modifier))
    o = dict(P=P, W=W, H=H)
    o["radii"] = np.zeros(P, np.int32)
    o["means2D"] = np.zeros((P, 2), dtype)
    o["depths"] = np.zeros(P, dtype)
    o["cov3D"] = np.zeros((P, 6), dtype)
    o["rgb"] = np.zeros((P, 3), dtype)
    o["conic_opacity"] = np.zeros((P, 4), dtype)
    o["clamped"] = np.zeros((P, 3), np.uint8)
    o["tiles_touched"] = np.zeros(P, np.uint32)
    o["point_offsets"] = np.zeros(P, np.uint32)
    R = getattr(L, "orc_preprocess_" + suf)(
        ctypes.c_int(P), ctypes.c_int(sh_degree), ctypes.c_int(M), _p(means3D), _p(scales), real(scale_modifier),
        _p(rotations), _p(opacities), _p(shs), _p(cov3D_precomp), _p(colors_precomp), _p(viewmatrix), _p(projmatrix),
        _p(campos), ctypes.c_int(W), ctypes.c_int(H), real(tanfovx), real(tanfovy), _p(o["radii"]), _p(o["means2D"]),
        _p(o["depths"]), _p(o["cov3D"]), _p(o["rgb"]), _p(o["conic_opacity"]), _p(o["clamped"]),
        _p(o["tiles_touched"]), _p(o["point_offsets"]))
    R = int(R)
    o["num_rendered"] = R
    gx, gy = (W + 15) // 16, (H + 15) // 16
    T = gx * gy
    o["keys_unsorted"] = np.zeros(R, np.uint64)
    o["values_unsorted"] = np.zeros(R, np.uint32)
    o["keys"] = np.zeros(R, np.uint64)
    o["point_list"] = np.zeros(R, np.uint32)
    o["ranges"] = np.zeros((T, 2), np.uint32)
    getattr(L, "orc_binning_" + suf)(
        ctypes.c_int(P), ctypes.c_int(W), ctypes.c_int(H), _p(o["radii"]), _p(o["means2D"]), _p(o["depths"]),
        _p(o["point_offsets"]), ctypes.c_int64(R), _p(o["keys_unsorted"]), _p(o["values_unsorted"]), _p(o["keys"]),
        _p(o["point_list"]), _p(o["ranges"]))
    colors = colors_precomp if colors_precomp is not None else o["rgb"]
    o["final_T"] = np.zeros((H, W), dtype)
    o["n_contrib"] = np.zeros((H, W), np.uint32)
    o["color"] = np.zeros((3, H, W), dtype)
    o["feature"] = np.zeros((3, H, W), dtype) if include_feature else np.zeros(1, dtype)
    kex, kco = ctypes.c_int64(0), ctypes.c_int64(0)
    getattr(L, "orc_render_" + suf)(
        ctypes.c_int(W), ctypes.c_int(H), _p(o["ranges"]), _p(o["point_list"]), _p(o["means2D"]), _p(colors),
        _p(language_feature if include_feature else None), _p(o["conic_opacity"]), _p(bg),
        ctypes.c_int(1 if include_feature else 0), _p(o["final_T"]), _p(o["n_contrib"]), _p(o["color"]),
        _p(o["feature"]), ctypes.byref(kex), ctypes.byref(kco))
    o["k_exam"], o["k_contrib"] = int(kex.value), int(kco.value)
    o["_inputs"] = dict(means3D=means3D, opacities=opacities, shs=shs, colors_precomp=colors_precomp,
                        language_feature=language_feature, scales=scales, rotations=rotations,
                        cov3D_precomp=cov3D_precomp, viewmatrix=viewmatrix, projmatrix=projmatrix, campos=campos,
                        bg=bg, tanfovx=tanfovx, tanfovy=tanfovy, scale_modifier=scale_modifier, sh_degree=sh_degree,
                        include_feature=include_feature, M=M, dtype=dtype)
    return o


def backward(fw: Dict[str, np.ndarray], dL_dcolor, dL_dfeature: Optional[np.ndarray]) -> Dict[str, np.ndarray]:
    L = lib()
    i = fw["_inputs"]
    dtype = i["dtype"]
    suf = "f32" if dtype == np.float32 else "f64"
    real = ctypes.c_float if dtype == np.float32 else ctypes.c_double
    P, W, H, M = fw["P"], fw["W"], fw["H"], i["M"]
    inc = i["include_feature"]
    dL_dcolor = _np(dL_dcolor, dtype)
    dL_dfeature = _np(dL_dfeature, dtype) if inc else None
    g = dict()
    g["means2D"] = np.zeros((P, 3), dtype)
    g["conic"] = np.zeros((P, 4), dtype)
    g["opacities"] = np.zeros((P, 1), dtype)
    g["colors"] = np.zeros((P, 3), dtype)
    g["language_feature"] = np.zeros((P, 3), dtype) if inc else np.zeros(1, dtype)
    g["means3D"] = np.zeros((P, 3), dtype)
    g["cov3D"] = np.zeros((P, 6), dtype)
    g["shs"] = np.zeros((P, M, 3), dtype)
    g["scales"] = np.zeros((P, 3), dtype)
    g["rotations"] = np.zeros((P, 4), dtype)
    colors = i["colors_precomp"] if i["colors_precomp"] is not None else fw["rgb"]
    getattr(L, "orc_render_bw_" + suf)(
        ctypes.c_int(P), ctypes.c_int(W), ctypes.c_int(H), _p(fw["ranges"]), _p(fw["point_list"]), _p(i["bg"]),
        _p(fw["means2D"]), _p(fw["conic_opacity"]), _p(colors), _p(i["language_feature"] if inc else None),
        _p(fw["final_T"]), _p(fw["n_contrib"]), _p(dL_dcolor), _p(dL_dfeature), ctypes.c_int(1 if inc else 0),
        _p(g["means2D"]), _p(g["conic"]), _p(g["opacities"]), _p(g["colors"]), _p(g["language_feature"]))
    cov3Ds = i["cov3D_precomp"] if i["cov3D_precomp"] is not None else fw["cov3D"]
    getattr(L, "orc_preprocess_bw_" + suf)(
        ctypes.c_int(P), ctypes.c_int(i["sh_degree"]), ctypes.c_int(M), _p(i["means3D"]), _p(fw["radii"]),
        _p(i["shs"]), _p(fw["clamped"]), _p(i["scales"]), _p(i["rotations"]), real(i["scale_modifier"]), _p(cov3Ds),
        _p(i["viewmatrix"]), _p(i["projmatrix"]), ctypes.c_int(W), ctypes.c_int(H), real(i["tanfovx"]),
        real(i["tanfovy"]), _p(i["campos"]), _p(g["means2D"]), _p(g["conic"]), _p(g["means3D"]), _p(g["colors"]),
        _p(g["cov3D"]), _p(g["shs"]), _p(g["scales"]), _p(g["rotations"]))
    return g


def mark_visible(means3D, viewmatrix, dtype=np.float32) -> np.ndarray:
    means3D = _np(means3D, dtype)
    viewmatrix = _np(viewmatrix, dtype).reshape(-1)
    out = np.zeros(means3D.shape[0], np.uint8)
    getattr(lib(), "orc_mark_visible_" + ("f32" if dtype == np.float32 else "f64"))(
        ctypes.c_int(means3D.shape[0]), _p(means3D), _p(viewmatrix), _p(out))
    return out.astype(bool)


# ---- L1 loss on the rendered images -------------------------------------------------------------
def l1_loss(network_output, gt, mask=None) -> np.float32:
    """utils/loss_utils.py:17-18 `torch.abs((network_output - gt)).mean()`, with train.py:96-99's masking
    (`l1_loss(x * mask, gt * mask)`, mask broadcast over the channel dimension) when `mask` is given.
    float32 products and difference as PyTorch computes them; the mean is accumulated in float64."""
    x = np.asarray(network_output, np.float32)
    g = np.asarray(gt, np.float32)
    if mask is not None:
        m = np.broadcast_to(np.asarray(mask, np.float32), x.shape)
        x, g = x * m, g * m
    d = (x - g).astype(np.float32)
    return np.float32(np.abs(d).astype(np.float64).mean()) if d.size else np.float32("nan")


def l1_loss_grad(network_output, gt, mask=None, grad_out: float = 1.0) -> np.ndarray:
    """Gradient of `l1_loss` w.r.t. network_output: grad_out * sign(x*m - gt*m) * m / numel (sign(0) = 0)."""
    x = np.asarray(network_output, np.float32)
    g = np.asarray(gt, np.float32)
    m = np.ones_like(x) if mask is None else np.broadcast_to(np.asarray(mask, np.float32), x.shape)
    d = (x * m - g * m).astype(np.float32)
    scale = np.float32(np.float32(grad_out) * np.float32(1.0 / max(d.size, 1)))
    return (scale * np.sign(d) * m).astype(np.float32)


# ---- activations in front of the rasterizer -----------------------------------------------------
def activate(scaling=None, rotation=None, opacity=None, language_feature=None):
    """(exp(scaling), normalize(rotation), sigmoid(opacity), f / (||f|| + 1e-9)) in float32:
    scene/gaussian_model.py:33,37,40 (torch.exp, torch.sigmoid, F.normalize with eps 1e-12) as read by
    get_scaling / get_rotation / get_opacity (:134-155), and gaussian_renderer/__init__.py:86-88.  None -> None."""
    f32 = np.float32
    out = [None, None, None, None]
    if scaling is not None:
        out[0] = np.exp(np.asarray(scaling, f32)).astype(f32)
    if rotation is not None:
        q = np.asarray(rotation, f32)
        n = np.sqrt((q * q).sum(axis=1, keepdims=True, dtype=f32)).astype(f32)
        out[1] = (q / np.maximum(n, f32(1e-12))).astype(f32)
    if opacity is not None:
        x = np.asarray(opacity, f32)
        with np.errstate(over="ignore"):  # exp(-x) -> inf for very negative x, sigmoid -> 0 as in PyTorch
            out[2] = (f32(1) / (f32(1) + np.exp(-x).astype(f32))).astype(f32)
    if language_feature is not None:
        f = np.asarray(language_feature, f32)
        n = np.sqrt((f * f).sum(axis=1, keepdims=True, dtype=f32)).astype(f32)
        out[3] = (f / (n + f32(1e-9))).astype(f32)
    return tuple(out)


def activate_grad(scaling=None, rotation=None, opacity=None, language_feature=None, g_scales=None, g_rotations=None,
                  g_opacities=None, g_feature=None):
    """Gradients of `activate` w.r.t. the raw parameters (float64 inside, float32 out), as autograd derives them from
    the reference expressions: clamp_min passes the gradient where ||q|| >= eps; norm's backward is 0 at ||f|| = 0."""
    out = [None, None, None, None]
    if scaling is not None and g_scales is not None:
        out[0] = (np.asarray(g_scales, np.float64) * np.exp(np.asarray(scaling, np.float64))).astype(np.float32)
    if rotation is not None and g_rotations is not None:
        q, g = np.asarray(rotation, np.float64), np.asarray(g_rotations, np.float64)
        n = np.sqrt((q * q).sum(axis=1, keepdims=True))
        d = np.maximum(n, 1e-12)
        k = np.where(n >= 1e-12, (g * q).sum(axis=1, keepdims=True) / (d * d * np.where(n > 0, n, 1.0)), 0.0)
        out[1] = (g / d - q * k).astype(np.float32)
    if opacity is not None and g_opacities is not None:
        y = 1.0 / (1.0 + np.exp(-np.asarray(opacity, np.float64)))
        out[2] = (np.asarray(g_opacities, np.float64) * y * (1.0 - y)).astype(np.float32)
    if language_feature is not None and g_feature is not None:
        f, g = np.asarray(language_feature, np.float64), np.asarray(g_feature, np.float64)
        n = np.sqrt((f * f).sum(axis=1, keepdims=True))
        d = n + 1e-9
        k = np.where(n > 0, (g * f).sum(axis=1, keepdims=True) / (d * d * np.where(n > 0, n, 1.0)), 0.0)
        out[3] = (g / d - f * k).astype(np.float32)
    return tuple(out)


# ---- ground-truth language-feature map ----------------------------------------------------------
def language_feature_map(seg_map, feature_map, feature_level, image_height, image_width):
    """scene/cameras.py:58-92: seg = seg_map[:, y, x].long() over the image's meshgrid (the top-left H x W of the map),
    mask = seg != -1, point_feature = feature_map[seg[level]] as [C,H,W] (numpy's negative index wraps like torch's:
    -1 selects the last row).  Returns (point_feature float32 [C,H,W], mask bool [1,H,W])."""
    if feature_level not in (0, 1, 2, 3):
        raise ValueError("feature_level=", feature_level)
    seg = np.asarray(seg_map)[feature_level, :image_height, :image_width]
    seg = np.trunc(seg).astype(np.int64)  # .long() truncates toward zero
    mask = (seg != -1)[None]
    feat = np.asarray(feature_map, np.float32)[seg]  # [H,W,C]; raises IndexError outside [-K, K)
    return np.ascontiguousarray(feat.transpose(2, 0, 1)), mask


# ---- Adam step ----------------------------------------------------------------------------------
def adam_step(param, grad, exp_avg, exp_avg_sq, step, lr, beta1=0.9, beta2=0.999, eps=1e-15, visible=None):
    """One torch.optim.Adam step (torch/optim/adam.py _single_tensor_adam: amsgrad off, no weight decay) as the
    reference configures it (scene/gaussian_model.py:227: Adam(l, lr=0.0, eps=1e-15), per-group lr), in float64.
    `step` is the count after the increment.  visible [P] bool: rows not visible keep param and moments.
    Returns (param, exp_avg, exp_avg_sq) as float64 arrays."""
    p, g = np.asarray(param, np.float64), np.asarray(grad, np.float64)
    m0, v0 = np.asarray(exp_avg, np.float64), np.asarray(exp_avg_sq, np.float64)
    m = m0 + (g - m0) * (1.0 - beta1)
    v = v0 * beta2 + (1.0 - beta2) * g * g
    bc1, bc2 = 1.0 - beta1 ** step, 1.0 - beta2 ** step
    denom = np.sqrt(v) / np.sqrt(bc2) + eps
    new = p - (lr / bc1) * (m / denom)
    if visible is not None:
        on = np.asarray(visible, bool).reshape((-1,) + (1,) * (p.ndim - 1))
        new, m, v = np.where(on, new, p), np.where(on, m, m0), np.where(on, v, v0)
    return new, m, v


def _round_operand(a, operand):
    """Operand rounding of the tensor-core variants of the CUDA decode kernel: None = none; "tf32" = fp32 rounded to the
    nearest value with 10 mantissa bits (what csrc/decoder.cu's tf32_rn does before the tensor core truncates);
    "tf32_trunc" = plain truncation (raw fp32 operands fed to kind::tf32); "f16" = round to nearest fp16, clamped."""
    if operand is None:
        return a
    a32 = np.asarray(a, np.float32)
    if operand == "f16":
        return np.clip(a32, -65504.0, 65504.0).astype(np.float16).astype(np.float64)
    u = a32.view(np.uint32)
    if operand == "tf32":
        u = (u + np.uint32(0x1000)) & np.uint32(0xFFFFE000)
    elif operand == "tf32_trunc":
        u = u & np.uint32(0xFFFFE000)
    else:
        raise ValueError(operand)
    return u.view(np.float32).astype(np.float64)


def _bf16_split(a):
    """(hi, lo) of the CUDA kernel's split-bf16 operands: hi = bf16(x), lo = bf16(x - hi), round to nearest even."""
    def rn(v32):
        u = v32.view(np.uint32)
        u = (u + np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1))) & np.uint32(0xFFFF0000)
        return u.view(np.float32)
    a32 = np.ascontiguousarray(a, np.float32)
    hi = rn(a32)
    lo = rn((a32 - hi).astype(np.float32))
    return hi.astype(np.float64), lo.astype(np.float64)


def decode_features(x, weights, biases, dtype=np.float64, operand=None):
    """Autoencoder.decode (autoencoder/model.py:42-46, decoder built at :18-25): Linear, then (ReLU, Linear) for every
    further layer, then x / x.norm(dim=-1, keepdim=True).  weights[l] is nn.Linear's [out, in] matrix.  float64 by
    default (the checker); dtype=np.float32 for the timed CPU baseline.  `operand` models the operand rounding of the
    CUDA kernel's tensor-core variants (activations and weights rounded before every product, fp32 accumulator stored
    between layers), which is where their tolerance comes from."""
    h = _np(x, dtype)
    for l, (w, b) in enumerate(zip(weights, biases)):
        if l > 0:
            h = np.maximum(h, 0)
        if operand is None:
            h = h @ _np(w, dtype).T + _np(b, dtype)
        elif operand == "bf16x2":  # three bf16 MMAs per product: hi*hi + lo*hi + hi*lo, fp32 accumulator
            (hh, hl), (wh, wl) = _bf16_split(h), _bf16_split(w)
            h = (hh @ wh.T + hl @ wh.T + hh @ wl.T).astype(np.float32).astype(np.float64)
            h = (h.astype(np.float32) + _np(b, np.float32)).astype(np.float64)
        else:
            h = (_round_operand(h, operand) @ _round_operand(w, operand).T).astype(np.float32).astype(np.float64)
            h = (h.astype(np.float32) + _np(b, np.float32)).astype(np.float64)
    return h / np.sqrt((h * h).sum(-1, keepdims=True))


def relevancy(embed, phrases, n_pos, dtype=np.float64):
    """OpenCLIPNetwork.get_relevancy for every positive phrase (eval/openclip_encoder.py:42-56): sims = embed @ p.T,
    softmax over the pair (10 * positive, 10 * negative_j) for every negative j, the pair whose positive probability is
    lowest is returned.  Returns [n_pos, N, 2] (column 0 is what get_max_across keeps, :96-111)."""
    e, p = _np(embed, dtype), _np(phrases, dtype)
    sims = e @ p.T
    neg = sims[:, n_pos:]
    out = np.empty((n_pos, e.shape[0], 2), dtype)
    for j in range(n_pos):
        pos = np.repeat(sims[:, j:j + 1], neg.shape[1], 1)
        pair = 10 * np.stack((pos, neg), -1)
        pair = np.exp(pair - pair.max(-1, keepdims=True))
        sm = pair / pair.sum(-1, keepdims=True)
        best = sm[..., 0].argmin(1)
        out[j] = sm[np.arange(e.shape[0]), best]
    return out
