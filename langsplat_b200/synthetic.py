"""Seeded synthetic scenes, cameras and upstream gradients for tests and bench.py.

Implements SURVEY.md Appendix C.  Camera matrices follow the reference's conventions:
`viewmatrix` = (world->view)^T, `projmatrix` = viewmatrix @ P^T with P from
utils/graphics_utils.py:51-71 (znear 0.01, zfar 100), exactly what scene/cameras.py:54-57 builds.
Everything is generated on the CPU in float32 from a seeded torch.Generator so every box sees the
same data; callers upload with .to(device).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict

import torch

# (P, W, H) of BASELINE.json's configs; C4 uses C2's resolution with 3M Gaussians and 8 cameras.
CONFIGS = {
    1: dict(P=16384, W=256, H=256),
    2: dict(P=1_000_000, W=1920, H=1080),
    3: dict(P=500_000, W=1440, H=1080),
    4: dict(P=3_000_000, W=1920, H=1080),
    5: dict(P=6_000_000, W=3840, H=2160),
}


@dataclass
class Camera:
    image_width: int
    image_height: int
    tanfovx: float
    tanfovy: float
    viewmatrix: torch.Tensor  # [4,4] f32, transposed world->view (flat index m[0]*x+m[4]*y+m[8]*z+m[12])
    projmatrix: torch.Tensor  # [4,4] f32, full projection, same convention
    campos: torch.Tensor  # [3]

    def to(self, device):
        return Camera(self.image_width, self.image_height, self.tanfovx, self.tanfovy,
                      self.viewmatrix.to(device), self.projmatrix.to(device), self.campos.to(device))


def make_gaussians(P: int, seed: int, sh_coeffs: int = 16) -> Dict[str, torch.Tensor]:
    g = torch.Generator().manual_seed(seed)
    means3D = torch.rand(P, 3, generator=g) * 2.0 - 1.0
    mu = math.log(0.5 * max(P, 1) ** (-1.0 / 3.0))
    scales = torch.exp(torch.randn(P, 3, generator=g) * 0.6 + mu)
    rotations = torch.nn.functional.normalize(torch.randn(P, 4, generator=g), dim=-1)
    opacities = torch.sigmoid(torch.randn(P, 1, generator=g) * 2.0)
    shs = torch.randn(P, sh_coeffs, 3, generator=g) * 0.1
    shs[:, 0, :] = torch.randn(P, 3, generator=g)
    lang = torch.nn.functional.normalize(torch.randn(P, 3, generator=g), dim=-1)
    return dict(means3D=means3D.contiguous(), scales=scales.contiguous(), rotations=rotations.contiguous(),
                opacities=opacities.contiguous(), shs=shs.contiguous(), language_feature=lang.contiguous())


def projection_matrix(znear: float, zfar: float, tanx: float, tany: float) -> torch.Tensor:
    """utils/graphics_utils.py:51-71 with tanHalfFov given directly."""
    top = tany * znear
    right = tanx * znear
    Pm = torch.zeros(4, 4, dtype=torch.float64)
    Pm[0, 0] = 2.0 * znear / (2 * right)
    Pm[1, 1] = 2.0 * znear / (2 * top)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    return Pm


def make_camera(W: int, H: int, k: int, seed: int, radius: float = 2.5) -> Camera:
    g = torch.Generator().manual_seed(seed * 1000 + k)
    jit = (torch.rand(2, generator=g, dtype=torch.float64) * 2 - 1)
    theta = 2 * math.pi * k / 8 + 0.1 + 0.05 * float(jit[0])
    height = 0.1 * float(jit[1])
    c = torch.tensor([radius * math.sin(theta), height, -radius * math.cos(theta)], dtype=torch.float64)
    f = -c / c.norm()
    down = torch.tensor([0.0, 1.0, 0.0], dtype=torch.float64)
    r = torch.linalg.cross(down, f)
    r = r / r.norm()
    d2 = torch.linalg.cross(f, r)
    W2C = torch.stack([r, d2, f])  # rows
    t = -W2C @ c
    V = torch.eye(4, dtype=torch.float64)
    V[:3, :3] = W2C
    V[:3, 3] = t
    fx = fy = 0.9 * W
    tanx = W / (2 * fx)
    tany = H / (2 * fy)
    Pp = projection_matrix(0.01, 100.0, tanx, tany)
    viewmatrix = V.t().to(torch.float32).contiguous()
    projmatrix = (viewmatrix.to(torch.float64) @ Pp.t()).to(torch.float32).contiguous()
    return Camera(W, H, float(tanx), float(tany), viewmatrix, projmatrix, c.to(torch.float32))


def make_grads(W: int, H: int, seed: int, feature_only: bool = False):
    g = torch.Generator().manual_seed(seed * 7919 + 17)
    dcol = torch.rand(3, H, W, generator=g) * 2 - 1
    dfeat = torch.rand(3, H, W, generator=g) * 2 - 1
    if feature_only:
        mask = (torch.rand(1, H, W, generator=g) < 0.8).to(torch.float32)
        dcol = torch.zeros_like(dcol)
        dfeat = dfeat * mask
    return dcol.contiguous(), dfeat.contiguous()


def make_config(config: int, P: int | None = None, W: int | None = None, H: int | None = None, n_cameras: int = 1):
    """Scene + cameras for one of BASELINE.json's configs (optionally down-scaled for tests)."""
    c = CONFIGS[config]
    P = c["P"] if P is None else P
    W = c["W"] if W is None else W
    H = c["H"] if H is None else H
    scene = make_gaussians(P, 1000 + config)
    cams = [make_camera(W, H, k, 2000 + config) for k in range(n_cameras)]
    return scene, cams
