"""Deterministic synthetic scenes and reference-convention cameras.

There are no datasets or trained models offline, so every benchmark / parity case uses a scene
generated here on the CPU with a seeded `torch.Generator` (identical bits for oracle and GPU).

Camera matrices follow the reference exactly:
  getWorld2View2        /root/reference/utils/graphics_utils.py:49-87
  getProjectionMatrix   /root/reference/utils/graphics_utils.py:97-133
  transposes / product  /root/reference/scene/cameras.py:103-121
"""
from __future__ import annotations

import math
from typing import NamedTuple, Optional

import numpy as np
import torch


class Camera(NamedTuple):
    image_height: int
    image_width: int
    FoVx: float
    FoVy: float
    world_view_transform: torch.Tensor  # [4,4] = W2C^T
    full_proj_transform: torch.Tensor   # [4,4] = W2C^T @ P^T
    camera_center: torch.Tensor         # [3]

    @property
    def tanfovx(self) -> float:
        return math.tan(self.FoVx * 0.5)

    @property
    def tanfovy(self) -> float:
        return math.tan(self.FoVy * 0.5)

    def to(self, device) -> "Camera":
        return self._replace(world_view_transform=self.world_view_transform.to(device),
                             full_proj_transform=self.full_proj_transform.to(device),
                             camera_center=self.camera_center.to(device))


def _world2view(R: np.ndarray, t: np.ndarray) -> np.ndarray:
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.transpose()
    Rt[:3, 3] = t
    Rt[3, 3] = 1.0
    C2W = np.linalg.inv(Rt)
    Rt = np.linalg.inv(C2W)       # translate = 0, scale = 1 (reference defaults)
    return np.float32(Rt)


def _projection(znear: float, zfar: float, fovX: float, fovY: float) -> torch.Tensor:
    tanY, tanX = math.tan(fovY / 2), math.tan(fovX / 2)
    top, right = tanY * znear, tanX * znear
    bottom, left = -top, -right
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def make_camera(width: int, height: int, yaw: float = 0.0, center=(0.0, -0.6, -4.0),
                focal_frac: float = 0.75) -> Camera:
    """Camera on an orbit about the y axis.  yaw = 0: at `center`, looking down +z (R = I).
    fx = fy = focal_frac * width  =>  tanfovx = 1 / (2 focal_frac), tanfovy = height / (2 fx)."""
    f = focal_frac * width
    FoVx = 2.0 * math.atan(width / (2.0 * f))
    FoVy = 2.0 * math.atan(height / (2.0 * f))
    c, s = math.cos(yaw), math.sin(yaw)
    Ry = np.array([[c, 0.0, s], [0.0, 1.0, 0.0], [-s, 0.0, c]])   # camera-to-world rotation
    C = Ry @ np.asarray(center, dtype=np.float64)
    t = -Ry.T @ C
    wvt = torch.tensor(_world2view(Ry, t)).transpose(0, 1).contiguous()
    proj = _projection(0.01, 100.0, FoVx, FoVy).transpose(0, 1)
    full = wvt.unsqueeze(0).bmm(proj.unsqueeze(0)).squeeze(0).contiguous()
    center_t = wvt.inverse()[3, :3].contiguous()
    return Camera(int(height), int(width), FoVx, FoVy, wvt, full, center_t)


class Scene(NamedTuple):
    means3D: torch.Tensor    # [P,3]
    scales: torch.Tensor     # [P,3] (already exp-activated)
    rotations: torch.Tensor  # [P,4] unit quaternions (w,x,y,z)
    opacities: torch.Tensor  # [P,1] in (0,1)
    shs: torch.Tensor        # [P,M,3]
    sh_degree: int

    def to(self, device) -> "Scene":
        return Scene(*[t.to(device) if isinstance(t, torch.Tensor) else t for t in self])


PRESETS = {
    # opacity = sigmoid(N(0, 2^2)) -- very opaque splats: a view blends ~7 % of its Gaussians
    "opaque": (0.0, 2.0),
    # a second operating point: opacity = sigmoid(N(-2, 1)), median 0.12 -- light passes through many
    # more layers, several times more Gaussians receive a gradient (the regime of a partly trained scene)
    "translucent": (-2.0, 1.0),
}


def make_scene(P: int, sh_degree: int = 3, seed: int = 0, s_med: Optional[float] = None, preset: str = "opaque") -> Scene:
    """70 % foreground blob N(0, diag(1.2, 0.45, 1.2)^2), 30 % background shell r ~ U(6,16);
    log-scales ~ N(ln s_med, 0.7^2) (background x r/3); s_med(P) = 0.006 (3e6/P)^(1/3); opacity per PRESETS."""
    g = torch.Generator().manual_seed(seed)
    M = (sh_degree + 1) ** 2
    if s_med is None:
        s_med = 0.006 * (3.0e6 / max(P, 1)) ** (1.0 / 3.0)
    n_bg = int(round(0.3 * P))
    n_fg = P - n_bg
    fg = torch.randn(n_fg, 3, generator=g) * torch.tensor([1.2, 0.45, 1.2])
    d = torch.randn(n_bg, 3, generator=g)
    d = d / d.norm(dim=1, keepdim=True).clamp_min(1e-12)
    r = 6.0 + 10.0 * torch.rand(n_bg, 1, generator=g)
    means = torch.cat([fg, d * r], dim=0)
    log_s = math.log(s_med) + 0.7 * torch.randn(P, 3, generator=g)
    scales = torch.exp(log_s)
    scales[n_fg:] = scales[n_fg:] * (r / 3.0)
    q = torch.randn(P, 4, generator=g)
    q = q / q.norm(dim=1, keepdim=True).clamp_min(1e-12)
    o_mean, o_std = PRESETS[preset]
    opac = torch.sigmoid(o_mean + o_std * torch.randn(P, 1, generator=g))
    shs = torch.empty(P, M, 3)
    shs[:, 0, :] = 0.5 * torch.randn(P, 3, generator=g)
    if M > 1:
        shs[:, 1:, :] = 0.1 * torch.randn(P, M - 1, 3, generator=g)
    # interleave foreground / background so index order carries no spatial structure
    perm = torch.randperm(P, generator=g)
    return Scene(means[perm].contiguous(), scales[perm].contiguous(), q[perm].contiguous(),
                 opac[perm].contiguous(), shs[perm].contiguous(), sh_degree)


def make_grad_image(width: int, height: int, seed: int = 1) -> torch.Tensor:
    """Fixed dL/d(colour) ~ N(0,1) used as the upstream gradient (decouples the op from any loss)."""
    g = torch.Generator().manual_seed(seed)
    return torch.randn(3, height, width, generator=g)
