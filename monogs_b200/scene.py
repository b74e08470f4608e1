"""Seeded synthetic scenes and cameras for the parity tests and benchmarks.

Everything is generated on the CPU with torch.Generator().manual_seed(seed) in float32 so that the CPU
oracle, the GPU oracle and the product see identical bits.  Camera conventions follow the MonoGS tree:
getProjectionMatrix2 (gaussian_splatting/utils/graphics_utils.py:56-77), the Camera properties
(utils/camera_utils.py:94-108) and the settings built in gaussian_renderer/__init__.py:61-75.
"""
import math
from typing import NamedTuple, Optional

import torch

CONFIGS = {
    # name: (P, W, H, fx, fy, cx, cy, sh_degree)
    "C1": (50_000, 640, 480, 535.4, 539.2, 320.1, 247.6, 0),
    "C2": (200_000, 640, 480, 535.4, 539.2, 320.1, 247.6, 0),
    "C3": (500_000, 1200, 680, 600.0, 600.0, 599.5, 339.5, 0),
    "C5": (2_000_000, 1920, 1080, 1000.0, 1000.0, 959.5, 539.5, 3),
}


class Camera(NamedTuple):
    W: int
    H: int
    fx: float
    fy: float
    cx: float
    cy: float
    tanfovx: float
    tanfovy: float
    viewmatrix: torch.Tensor       # [4,4] = T_w2c transposed (memory = column-major T)
    projmatrix: torch.Tensor       # [4,4] = viewmatrix @ projection
    projmatrix_raw: torch.Tensor   # [4,4] = projection (already transposed)
    campos: torch.Tensor           # MonoGS hands the whole view matrix here


def projection_matrix(znear, zfar, cx, cy, fx, fy, W, H):
    """OpenGL-style intrinsics matrix of getProjectionMatrix2, returned un-transposed."""
    left = ((2 * cx - W) / W - 1.0) * W / 2.0
    right = ((2 * cx - W) / W + 1.0) * W / 2.0
    top = ((2 * cy - H) / H + 1.0) * H / 2.0
    bottom = ((2 * cy - H) / H - 1.0) * H / 2.0
    left, right = znear / fx * left, znear / fx * right
    top, bottom = znear / fy * top, znear / fy * bottom
    Pm = torch.zeros(4, 4)
    Pm[0, 0] = 2.0 * znear / (right - left)
    Pm[1, 1] = 2.0 * znear / (top - bottom)
    Pm[0, 2] = (right + left) / (right - left)
    Pm[1, 2] = (top + bottom) / (top - bottom)
    Pm[3, 2] = 1.0
    Pm[2, 2] = 1.0 * zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    return Pm


def se3_exp(tau: torch.Tensor) -> torch.Tensor:
    """4x4 Exp of tau = (rho, theta) (utils/pose_utils.py:26-74 restated in float64)."""
    tau = tau.double()
    rho, theta = tau[:3], tau[3:]
    Wm = torch.zeros(3, 3, dtype=torch.float64)
    Wm[0, 1], Wm[0, 2], Wm[1, 0], Wm[1, 2], Wm[2, 0], Wm[2, 1] = -theta[2], theta[1], theta[2], -theta[0], -theta[1], theta[0]
    a = float(theta.norm())
    I = torch.eye(3, dtype=torch.float64)
    if a < 1e-5:
        R = I + Wm + 0.5 * Wm @ Wm
        V = I + 0.5 * Wm + Wm @ Wm / 6.0
    else:
        R = I + math.sin(a) / a * Wm + (1 - math.cos(a)) / a**2 * Wm @ Wm
        V = I + (1 - math.cos(a)) / a**2 * Wm + (a - math.sin(a)) / a**3 * Wm @ Wm
    T = torch.eye(4, dtype=torch.float64)
    T[:3, :3] = R
    T[:3, 3] = V @ rho
    return T


def make_camera(W, H, fx, fy, cx, cy, T_w2c: Optional[torch.Tensor] = None) -> Camera:
    T = torch.eye(4) if T_w2c is None else T_w2c.float()
    proj = projection_matrix(0.01, 100.0, cx, cy, fx, fy, W, H).transpose(0, 1).contiguous()
    view = T.transpose(0, 1).contiguous()
    full = (view.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    return Camera(W, H, fx, fy, cx, cy, W / (2.0 * fx), H / (2.0 * fy), view, full, proj, view.clone())


class Scene(NamedTuple):
    means3D: torch.Tensor    # [P,3]
    scales: torch.Tensor     # [P,3] activated
    rotations: torch.Tensor  # [P,4] normalised (w,x,y,z)
    opacities: torch.Tensor  # [P,1] activated
    shs: torch.Tensor        # [P,M,3]
    sh_degree: int


def make_scene(P, W, H, fx, fy, cx, cy, sh_degree=0, seed=0) -> Scene:
    g = torch.Generator().manual_seed(seed)
    u = (torch.rand(P, generator=g) * 1.2 - 0.1) * W
    v = (torch.rand(P, generator=g) * 1.2 - 0.1) * H
    z = torch.exp(torch.rand(P, generator=g) * (math.log(6.0) - math.log(0.5)) + math.log(0.5))
    behind = torch.rand(P, generator=g) < 0.02
    z_b = torch.rand(P, generator=g) * 1.2 - 1.0
    z = torch.where(behind, z_b, z)
    means = torch.stack([(u - cx) / fx * z, (v - cy) / fy * z, z], dim=1)
    r_px = torch.exp(torch.randn(P, generator=g) * 0.6 + math.log(1.5))
    sigma = (z.abs() * r_px / fx).unsqueeze(1) * torch.exp(torch.randn(P, 3, generator=g) * 0.3)
    q = torch.randn(P, 4, generator=g)
    q = q / q.norm(dim=1, keepdim=True)
    opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 1.5)
    M = (sh_degree + 1) ** 2
    shs = torch.zeros(P, M, 3)
    shs[:, 0, :] = (torch.rand(P, 3, generator=g) - 0.5) / 0.28209479177387814
    if M > 1:
        shs[:, 1:, :] = torch.randn(P, M - 1, 3, generator=g) * 0.05
    return Scene(means.float().contiguous(), sigma.float().contiguous(), q.float().contiguous(),
                 opac.float().contiguous(), shs.float().contiguous(), sh_degree)


def make_config(name, seed=0, P=None):
    Pn, W, H, fx, fy, cx, cy, deg = CONFIGS[name]
    if P is not None:
        Pn = P
    return make_scene(Pn, W, H, fx, fy, cx, cy, deg, seed), make_camera(W, H, fx, fy, cx, cy)


def output_grads(W, H, seed=1):
    g = torch.Generator().manual_seed(seed)
    gc = torch.randn(3, H, W, generator=g) / (3 * H * W)
    gd = torch.randn(1, H, W, generator=g) / (H * W)
    return gc.contiguous(), gd.contiguous()


GT_TAU = (0.010, -0.005, 0.008, 0.004, -0.003, 0.002)
