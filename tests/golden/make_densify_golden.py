#!/usr/bin/env python
"""Generates tests/golden/densify_golden.npz from the REFERENCE's own GaussianModel.

Run in the authoring container only (needs /root/reference; the GPU box and CI never run this):
    python tests/golden/make_densify_golden.py
It imports /root/reference/scene/gaussian_model.py unmodified and drives its own methods
(training_setup, add_densification_stats, densify_and_prune, prune_points, reset_opacity) on CPU tensors.
Three things the container lacks are stood in for, none of them on the arithmetic being pinned:
  * `plyfile` and `simple_knn._C` (imported at module level, unused by these methods) are empty stub modules;
  * the hard-wired device="cuda" of torch.zeros/ones/eye (gaussian_model.py:506-508,511,519,548,
    general_utils.py:84) is redirected to the CPU;
  * `torch.normal(mean=means, std=stds)` (gaussian_model.py:521), the one random draw of the path, returns
    mean + std * Z[copy, source Gaussian] for a stored standard-normal table Z, so that the draw is an input.
"""
import os
import sys
import types
from argparse import Namespace

import numpy as np
import torch
from torch import nn

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"

# ---- stand-ins (see the docstring) --------------------------------------------------------------
ply = types.ModuleType("plyfile")
ply.PlyData = ply.PlyElement = object
sys.modules.setdefault("plyfile", ply)      # (tests may have installed gaussian_splatting_b200.plyfile_compat)
knn = types.ModuleType("simple_knn")
knn_c = types.ModuleType("simple_knn._C")
knn_c.distCUDA2 = None
knn._C = knn_c
sys.modules["simple_knn"] = knn
sys.modules["simple_knn._C"] = knn_c


def _cpu(fn):
    def wrapped(*a, **k):
        if str(k.get("device", "")).startswith("cuda"):
            k["device"] = "cpu"
        return fn(*a, **k)
    return wrapped


for _n in ("zeros", "ones", "eye", "tensor", "empty"):
    setattr(torch, _n, _cpu(getattr(torch, _n)))

sys.path.insert(0, REF)
from scene.gaussian_model import GaussianModel  # noqa: E402

GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
ATTR = {"xyz": "_xyz", "f_dc": "_features_dc", "f_rest": "_features_rest", "opacity": "_opacity",
        "scaling": "_scaling", "rotation": "_rotation"}
NOISE = {"Z": None, "model": None}


def _normal(mean=None, std=None, **_):
    """mean + std * Z[k, i] for the i-th selected Gaussian (rows of std are get_scaling[selected].repeat(2,1))."""
    gm, Z = NOISE["model"], NOISE["Z"]
    scal = gm.get_scaling.detach()
    ns = std.shape[0] // 2
    idx, j = [], 0
    for r in range(ns):
        while not torch.equal(scal[j], std[r]):
            j += 1
        idx.append(j)
        j += 1
    idx = torch.tensor(idx, dtype=torch.long)
    assert torch.equal(scal[idx].repeat(2, 1), std)
    z = torch.cat((Z[0][idx], Z[1][idx]), 0)
    return mean + std * z


torch.normal = _normal


def make_model(P, gen, sh_degree=3, log_scale_mean=-3.0):
    K = (sh_degree + 1) ** 2 - 1
    gm = GaussianModel(sh_degree)
    gm._xyz = nn.Parameter(torch.randn(P, 3, generator=gen) * 2)
    gm._features_dc = nn.Parameter(torch.randn(P, 1, 3, generator=gen) * 0.5)
    gm._features_rest = nn.Parameter(torch.randn(P, K, 3, generator=gen) * 0.1)
    gm._opacity = nn.Parameter(torch.randn(P, 1, generator=gen) * 2.5 - 2.0)
    gm._scaling = nn.Parameter(torch.randn(P, 3, generator=gen) * 0.9 + log_scale_mean)
    gm._rotation = nn.Parameter(torch.randn(P, 4, generator=gen))
    gm.max_radii2D = torch.zeros(P)
    gm.spatial_lr_scale = 1.0
    gm._exposure = nn.Parameter(torch.eye(3, 4)[None])
    gm.pretrained_exposures = None
    args = Namespace(percent_dense=0.01, position_lr_init=0.00016, position_lr_final=0.0000016,
                     position_lr_delay_mult=0.01, position_lr_max_steps=30000, feature_lr=0.0025, opacity_lr=0.025,
                     scaling_lr=0.005, rotation_lr=0.001, exposure_lr_init=0.01, exposure_lr_final=0.001,
                     exposure_lr_delay_steps=0, exposure_lr_delay_mult=0.0, iterations=30000)
    gm.training_setup(args)
    # a few optimizer steps so that both moments are non-trivial
    for _ in range(3):
        for g in GROUPS:
            p = getattr(gm, ATTR[g])
            p.grad = torch.randn(p.shape, generator=gen) * 0.01
        gm.optimizer.step()
        gm.optimizer.zero_grad(set_to_none=True)
    return gm


def dump(gm, prefix, out):
    for group in gm.optimizer.param_groups:
        p = group["params"][0]
        st = gm.optimizer.state[p]
        out[f"{prefix}{group['name']}"] = p.detach().numpy().copy()
        out[f"{prefix}m_{group['name']}"] = st["exp_avg"].numpy().copy()
        out[f"{prefix}v_{group['name']}"] = st["exp_avg_sq"].numpy().copy()
    out[f"{prefix}xyz_gradient_accum"] = gm.xyz_gradient_accum.numpy().copy()
    out[f"{prefix}denom"] = gm.denom.numpy().copy()
    out[f"{prefix}max_radii2D"] = gm.max_radii2D.numpy().copy()


def stats_steps(gm, gen, out, prefix, views=4):
    """train.py:212-215 for a few synthetic views; the per-view inputs are stored too."""
    P = gm.get_xyz.shape[0]
    for v in range(views):
        vis = torch.rand(P, generator=gen) < 0.7
        radii = (torch.rand(P, generator=gen) * 40).to(torch.int32) * vis
        vsp = types.SimpleNamespace(grad=torch.randn(P, 3, generator=gen) * 0.0004)
        out[f"{prefix}view{v}_vis"] = vis.numpy().copy()
        out[f"{prefix}view{v}_radii"] = radii.numpy().copy()
        out[f"{prefix}view{v}_grad"] = vsp.grad.numpy().copy()
        gm.max_radii2D[vis] = torch.max(gm.max_radii2D[vis], radii[vis])
        gm.add_densification_stats(vsp, vis)
    return radii


def case_densify(name, P, seed, max_grad, extent, size_threshold, out, log_scale_mean=-3.0):
    gen = torch.Generator().manual_seed(seed)
    gm = make_model(P, gen, log_scale_mean=log_scale_mean)
    radii = stats_steps(gm, gen, out, f"{name}_")
    dump(gm, f"{name}_in_", out)
    Z = torch.randn(2, 2 * P, 3, generator=gen)
    NOISE["Z"], NOISE["model"] = Z, gm
    out[f"{name}_noise"] = Z.numpy().copy()
    out[f"{name}_radii"] = radii.numpy().copy()
    out[f"{name}_args"] = np.array([max_grad, 0.005, extent, size_threshold if size_threshold else 0.0,
                                    gm.percent_dense], np.float64)
    gm.densify_and_prune(max_grad, 0.005, extent, size_threshold, radii)
    dump(gm, f"{name}_out_", out)
    print(name, "P", P, "->", gm.get_xyz.shape[0])
    return gm, gen


def main():
    out = {}
    # a: the training configuration (train.py:219-223): threshold 0.0002, size_threshold 20, extent 5
    gm, gen = case_densify("a", 400, 1, 0.0002, 5.0, 20, out)
    # follow-ups on the densified model: reset_opacity (:323-327) and a plain prune_points (:433-451)
    gm.reset_opacity()
    dump(gm, "a_reset_", out)
    mask = torch.rand(gm.get_xyz.shape[0], generator=gen) < 0.3
    gm.tmp_radii = torch.zeros(gm.get_xyz.shape[0])
    out["a_prune_mask"] = mask.numpy().copy()
    gm.prune_points(mask)
    dump(gm, "a_pruned_", out)
    # b: size_threshold None (iteration <= opacity_reset_interval, train.py:220): opacity-only pruning
    case_densify("b", 333, 2, 0.0002, 5.0, None, out)
    # c: nothing reaches the gradient threshold: only pruning happens
    case_densify("c", 257, 3, 1.0, 5.0, 20, out)
    # d: every Gaussian is large: splits only, and big_points_ws removes most of the result
    case_densify("d", 130, 4, 0.0001, 0.8, 20, out, log_scale_mean=-1.5)
    # e: a single Gaussian
    case_densify("e", 1, 5, 0.0, 5.0, 20, out)
    np.savez_compressed(os.path.join(HERE, "densify_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "densify_golden.npz"), len(out), "arrays")


if __name__ == "__main__":
    main()
