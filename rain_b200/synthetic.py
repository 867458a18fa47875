"""Seeded synthetic scenes and cameras for the parity tests and bench.py (BASELINE.md §3 / SURVEY.md §8(d)).

Everything is generated on the CPU with a seeded torch.Generator so that the oracle, the rebuilt reference CUDA and
this library see identical bytes.  Cameras are built the way the reference builds them
(scene/cameras.py:37-46, utils/graphics_utils.py:27-60): world_view_transform = W2V^T, full_proj = (P @ W2V)^T,
camera centre = inverse(world_view_transform)[3,:3].
"""
import math
from dataclasses import dataclass

import numpy as np
import torch

SCENE_RADIUS = 2.0


@dataclass
class Scene:
    means3D: torch.Tensor    # [P,3]
    scales: torch.Tensor     # [P,3]  (already activated: exp)
    rotations: torch.Tensor  # [P,4]  (already normalised)
    opacities: torch.Tensor  # [P,1]  (already activated: sigmoid)
    shs: torch.Tensor        # [P,16,3]
    sh_degree: int

    def to(self, device):
        return Scene(self.means3D.to(device), self.scales.to(device), self.rotations.to(device), self.opacities.to(device),
                     self.shs.to(device), self.sh_degree)


@dataclass
class Camera:
    width: int
    height: int
    tanfovx: float
    tanfovy: float
    viewmatrix: torch.Tensor   # [4,4] world_view_transform (transposed W2V)
    projmatrix: torch.Tensor   # [4,4] full_proj_transform
    campos: torch.Tensor       # [3]

    def to(self, device):
        return Camera(self.width, self.height, self.tanfovx, self.tanfovy, self.viewmatrix.to(device), self.projmatrix.to(device), self.campos.to(device))


def make_scene(P, sh_degree=3, seed=0, scale_mult=1.0):
    g = torch.Generator().manual_seed(seed)
    R = SCENE_RADIUS
    means = (torch.rand(P, 3, generator=g) * 2.0 - 1.0) * R
    s0 = 0.7 * R * max(P, 1) ** (-1.0 / 3.0) * scale_mult
    scales = torch.exp(torch.randn(P, 3, generator=g) * 0.6 + math.log(s0))
    q = torch.randn(P, 4, generator=g)
    rotations = q / q.norm(dim=1, keepdim=True)
    opacities = torch.sigmoid(torch.randn(P, 1, generator=g) * 2.0)
    shs = torch.randn(P, 16, 3, generator=g) * 0.2
    shs[:, 0, :] = torch.randn(P, 3, generator=g)
    return Scene(means.float().contiguous(), scales.float().contiguous(), rotations.float().contiguous(),
                 opacities.float().contiguous(), shs.float().contiguous(), sh_degree)


def _projection(znear, zfar, fovx, fovy):
    """getProjectionMatrix (utils/graphics_utils.py:40-60)"""
    t = math.tan(fovy / 2) * znear
    r = math.tan(fovx / 2) * znear
    Pm = torch.zeros(4, 4)
    Pm[0, 0] = 2.0 * znear / (2 * r)
    Pm[1, 1] = 2.0 * znear / (2 * t)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    return Pm


def make_camera(view_index, num_views, width, height, fovy_deg=50.0, radius=3.0 * SCENE_RADIUS, znear=0.01, zfar=100.0):
    # Fibonacci sphere position looking at the origin, up = +y
    k = view_index + 0.5
    n = max(num_views, 1)
    yv = 1.0 - 2.0 * k / n
    rad = math.sqrt(max(0.0, 1.0 - yv * yv))
    phi = k * math.pi * (3.0 - math.sqrt(5.0))
    c = np.array([math.cos(phi) * rad, yv, math.sin(phi) * rad], dtype=np.float64) * radius
    fwd = -c / np.linalg.norm(c)
    up = np.array([0.0, 1.0, 0.0])
    if abs(np.dot(fwd, up)) > 0.999:
        up = np.array([1.0, 0.0, 0.0])
    right = np.cross(up, fwd); right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    Rw2c = np.stack([right, down, fwd], axis=0)         # rows: camera x (right), y (down), z (forward)
    W2V = np.eye(4)
    W2V[:3, :3] = Rw2c
    W2V[:3, 3] = -Rw2c @ c
    fovy = math.radians(fovy_deg)
    fovx = 2.0 * math.atan(math.tan(fovy / 2) * width / height)
    wvt = torch.tensor(np.float32(W2V)).transpose(0, 1).contiguous()
    proj = _projection(znear, zfar, fovx, fovy).transpose(0, 1)
    full = wvt.unsqueeze(0).bmm(proj.unsqueeze(0)).squeeze(0).contiguous()
    campos = wvt.inverse()[3, :3].contiguous()
    return Camera(width, height, math.tan(fovx * 0.5), math.tan(fovy * 0.5), wvt, full, campos)


def make_grad(width, height, seed=1):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(3, height, width, generator=g).float().contiguous()


# the configs of BASELINE.json
CONFIGS = {
    "C1": dict(P=10_000, W=256, H=256, D=0),
    "C2": dict(P=300_000, W=800, H=800, D=3),
    "C3": dict(P=1_000_000, W=1920, H=1080, D=3),
    "C4": dict(P=6_000_000, W=1920, H=1080, D=3, views=64),
    "C5": dict(P=5_000_000),
}
