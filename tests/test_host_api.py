"""CPU: the host-side mirror of the reference's Python surface (RAST/diff_gaussian_rasterization/__init__.py) — names,
field order, argument checks and error messages — exercised without a GPU.  Anything that would reach a kernel must
fail loudly on CPU tensors: there is no CPU fallback."""
import inspect

import pytest
import torch

import diff_gaussian_rasterization as drop_in
import langsplat_b200 as pkg
from langsplat_b200 import rasterizer as R


def test_settings_field_order_matches_reference():
    assert R.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug", "include_feature")  # __init__.py:166-179


def test_forward_signature_matches_reference():
    sig = inspect.signature(R.GaussianRasterizer.forward)
    assert list(sig.parameters) == ["self", "means3D", "means2D", "opacities", "shs", "colors_precomp",
                                    "language_feature_precomp", "scales", "rotations", "cov3D_precomp"]  # :197
    assert all(sig.parameters[k].default is None for k in list(sig.parameters)[4:])
    assert list(inspect.signature(R.rasterize_gaussians).parameters) == [
        "means3D", "means2D", "sh", "colors_precomp", "language_feature_precomp", "opacities", "scales", "rotations",
        "cov3Ds_precomp", "raster_settings"]  # :21-32


def test_drop_in_package_exports_the_same_objects():
    assert drop_in.GaussianRasterizer is pkg.GaussianRasterizer
    assert drop_in.GaussianRasterizationSettings is pkg.GaussianRasterizationSettings
    assert drop_in.rasterize_gaussians is pkg.rasterize_gaussians
    assert hasattr(R.GaussianRasterizer, "markVisible")


def _rast():
    z = torch.zeros
    rs = R.GaussianRasterizationSettings(16, 16, 0.5, 0.5, z(3), 1.0, torch.eye(4), torch.eye(4), 3, z(3), False, False,
                                         True)
    return R.GaussianRasterizer(rs)


def test_exclusive_argument_errors_have_the_reference_wording():
    rast, P = _rast(), 4
    m, o = torch.zeros(P, 3), torch.ones(P, 1)
    sh, col = torch.zeros(P, 16, 3), torch.zeros(P, 3)
    sc, rot, cov = torch.ones(P, 3), torch.zeros(P, 4), torch.zeros(P, 6)
    with pytest.raises(Exception, match="Please provide excatly one of either SHs or precomputed colors!"):
        rast(m, m, o, scales=sc, rotations=rot)
    with pytest.raises(Exception, match="Please provide excatly one of either SHs or precomputed colors!"):
        rast(m, m, o, shs=sh, colors_precomp=col, scales=sc, rotations=rot)
    msg = "Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!"
    with pytest.raises(Exception, match=msg):
        rast(m, m, o, shs=sh)
    with pytest.raises(Exception, match=msg):
        rast(m, m, o, shs=sh, scales=sc)  # rotation missing
    with pytest.raises(Exception, match=msg):
        rast(m, m, o, shs=sh, scales=sc, rotations=rot, cov3D_precomp=cov)


def test_shape_check_and_no_cpu_path():
    rast, P = _rast(), 4
    o, sh = torch.ones(P, 1), torch.zeros(P, 16, 3)
    sc, rot = torch.ones(P, 3), torch.zeros(P, 4)
    with pytest.raises(RuntimeError, match=r"means3D must have dimensions \(num_points, 3\)"):
        rast(torch.zeros(P, 2), torch.zeros(P, 3), o, shs=sh, scales=sc, rotations=rot)
    with pytest.raises(R.LsrError, match="no CPU path"):
        rast(torch.zeros(P, 3), torch.zeros(P, 3), o, shs=sh, scales=sc, rotations=rot,
             language_feature_precomp=torch.zeros(P, 3))
    with pytest.raises(R.LsrError, match="no CPU path"):
        rast.markVisible(torch.zeros(P, 3))


def test_product_package_never_imports_the_oracle():
    import os
    root = os.path.dirname(os.path.abspath(pkg.__file__))
    for dirpath, _, files in os.walk(root):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("test oracle", ""), f"{f} mentions the oracle"
