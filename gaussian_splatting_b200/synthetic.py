"""Deterministic synthetic Gaussians and cameras for the benchmark configs.

Everything is generated on the CPU with a seeded torch.Generator so that every implementation
(this package, the CPU oracle, the reference CUDA build) sees identical bits.  Camera matrices
are built exactly as the reference does (scene/cameras.py:80-89, utils/graphics_utils.py:38-77):
`viewmatrix` = W2V transposed, `projmatrix` = (W2V^T @ Proj^T), `campos` = inverse(W2V^T)[3,:3].
"""
import math
from typing import NamedTuple

import numpy as np
import torch


class SynthConfig(NamedTuple):
    name: str
    P: int
    W: int
    H: int
    fovx: float
    antialiasing: bool
    seed: int


# BASELINE.json configs (C1..C4); C5 reuses C3's camera with varying P.
CONFIGS = {
    "C1": SynthConfig("C1", 100_000, 800, 800, 0.6911, False, 1),
    "C2": SynthConfig("C2", 1_000_000, 1297, 840, 1.19, False, 2),
    "C3": SynthConfig("C3", 3_000_000, 1920, 1080, 1.20, True, 3),
    "C4": SynthConfig("C4", 6_000_000, 3840, 2160, 1.20, False, 4),
}


def fov2focal(fov, pixels):
    return pixels / (2 * math.tan(fov / 2))


def focal2fov(focal, pixels):
    return 2 * math.atan(pixels / (2 * focal))


def get_world2view2(R, t):
    """utils/graphics_utils.py:38-49 with translate=0, scale=1."""
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.transpose()
    Rt[:3, 3] = t
    Rt[3, 3] = 1.0
    C2W = np.linalg.inv(Rt)
    Rt = np.linalg.inv(C2W)
    return np.float32(Rt)


def get_projection_matrix(znear, zfar, fovX, fovY):
    """utils/graphics_utils.py:51-71."""
    tanHalfFovY = math.tan(fovY / 2)
    tanHalfFovX = math.tan(fovX / 2)
    top = tanHalfFovY * znear
    bottom = -top
    right = tanHalfFovX * znear
    left = -right
    P = torch.zeros(4, 4)
    z_sign = 1.0
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = z_sign
    P[2, 2] = z_sign * zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def make_camera(W, H, fovx, yaw_deg=0.0, dist=None):
    """Camera looking at the origin from distance d = 1.3/tan(FoVy/2) + 1.3, rotated about y.

    Returns dict(image_width, image_height, tanfovx, tanfovy, viewmatrix, projmatrix, campos)
    of CPU float32 tensors / python floats (COLMAP convention: x right, y down, z forward).
    """
    fovy = focal2fov(fov2focal(fovx, W), H)
    d = 1.3 / math.tan(fovy / 2) + 1.3 if dist is None else dist
    a = math.radians(yaw_deg)
    # camera-to-world rotation (columns = camera axes in world); R passed to getWorld2View2 is C2W
    Rc2w = np.array([[math.cos(a), 0.0, math.sin(a)],
                     [0.0, 1.0, 0.0],
                     [-math.sin(a), 0.0, math.cos(a)]])
    center = Rc2w @ np.array([0.0, 0.0, -d])
    t = -Rc2w.T @ center                      # W2V translation
    w2v = torch.tensor(get_world2view2(Rc2w, t)).transpose(0, 1)
    proj = get_projection_matrix(0.01, 100.0, fovx, fovy).transpose(0, 1)
    full = (w2v.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0)
    campos = w2v.inverse()[3, :3]
    return dict(image_width=W, image_height=H, tanfovx=math.tan(fovx * 0.5), tanfovy=math.tan(fovy * 0.5),
                viewmatrix=w2v.contiguous(), projmatrix=full.contiguous(), campos=campos.contiguous())


def make_gaussians(P, W, H, seed, sh_degree=3, scale_mult=1.0):
    """Random Gaussians in the box of scene/dataset_readers.py:294 stretched to the aspect ratio.

    Returns CPU float32 tensors: means3D (P,3), scales (P,3) [post-activation], rotations (P,4)
    [normalised, (w,x,y,z)], opacities (P,1) [post-sigmoid], shs (P,(deg+1)^2,3).
    """
    g = torch.Generator().manual_seed(seed)
    a = W / H
    means = (torch.rand(P, 3, generator=g) * 2 - 1) * 1.3
    means[:, 0] *= a
    vol = (2.6 ** 3) * a
    s0 = 0.5 * (vol / P) ** (1.0 / 3.0) * scale_mult
    scales = torch.exp(math.log(s0) + 0.35 * torch.randn(P, 3, generator=g))
    rot = torch.nn.functional.normalize(torch.randn(P, 4, generator=g), dim=1)
    opac = torch.sigmoid(2.0 * torch.randn(P, 1, generator=g))
    M = (sh_degree + 1) ** 2
    shs = torch.randn(P, M, 3, generator=g) * 0.1
    shs[:, 0, :] = torch.randn(P, 3, generator=g) * 0.5
    return dict(means3D=means.contiguous(), scales=scales.contiguous(), rotations=rot.contiguous(),
                opacities=opac.contiguous(), shs=shs.contiguous())


def make_scene(cfg, P=None, yaw_deg=0.0):
    """(gaussians, camera) for a SynthConfig or config name; P overrides the Gaussian count."""
    if isinstance(cfg, str):
        cfg = CONFIGS[cfg]
    P = cfg.P if P is None else P
    gs = make_gaussians(P, cfg.W, cfg.H, cfg.seed)
    cam = make_camera(cfg.W, cfg.H, cfg.fovx, yaw_deg)
    cam["antialiasing"] = cfg.antialiasing
    return gs, cam


def make_targets(W, H, seed):
    g = torch.Generator().manual_seed(1000 + seed)
    return torch.rand(3, H, W, generator=g), torch.rand(1, H, W, generator=g)
