"""Shared test helpers: small seeded scenes and comparison metrics."""
import numpy as np
import torch

from gaussian_splatting_b200 import synthetic


def small_scene(P=300, W=48, H=40, seed=7, sh_degree=3, scale_mult=6.0, fovx=0.9, antialiasing=False,
                bg=(0.0, 0.0, 0.0), yaw_deg=0.0, behind_fraction=0.1):
    """Gaussians + settings dict (CPU tensors). A fraction of the Gaussians is pushed behind the
    near plane so that culling is exercised; sizes are inflated so tiles overlap heavily."""
    gs = synthetic.make_gaussians(P, W, H, seed, sh_degree=sh_degree, scale_mult=scale_mult)
    cam = synthetic.make_camera(W, H, fovx, yaw_deg)
    nb = int(P * behind_fraction)
    if nb:
        g = torch.Generator().manual_seed(seed + 99)
        idx = torch.randperm(P, generator=g)[:nb]
        gs["means3D"][idx] = gs["means3D"][idx] * 0.2 + cam["campos"][None, :] * 1.02
    s = dict(cam)
    s.update(bg=torch.tensor(bg, dtype=torch.float32), scale_modifier=1.0, sh_degree=sh_degree,
             antialiasing=antialiasing, prefiltered=False, debug=False)
    return gs, s


def np_inputs(gs):
    return {k: (v.numpy() if isinstance(v, torch.Tensor) else v) for k, v in gs.items()}


def np_settings(s):
    return {k: (v.numpy() if isinstance(v, torch.Tensor) else v) for k, v in s.items()}


def rel_err(a, b, eps=1e-12):
    """Norm-wise relative error  ||a-b||_inf / max(||b||_inf, eps)  (SURVEY.md §8c)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), eps)) if a.size else 0.0


def elem_rel_err(a, b, floor_frac=1e-3):
    """Max per-element relative error over elements with |b| > floor_frac * ||b||_inf."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    if not a.size:
        return 0.0
    m = np.abs(b) > floor_frac * np.abs(b).max()
    if not m.any():
        return 0.0
    return float((np.abs(a - b)[m] / np.abs(b)[m]).max())
