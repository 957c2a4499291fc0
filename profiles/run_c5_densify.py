#!/usr/bin/env python
"""BASELINE config C5 driven by the real density control instead of fresh synthetic sets:
one model at 1080p grows from 500K Gaussians to >= 5M through train steps (render, L1, backward, statistics,
SparseGaussianAdam) and gaussian_splatting_b200.densify.densify_and_prune every few steps.  Per growth round:
P before / after, the fwd+bwd time at that size, the densify_and_prune call time (CUDA events around the call,
its one host wait included), num_rendered, scratch arena size and regrow count.  The gradient threshold of each round
is the 75 % quantile of the accumulated statistics (the run is a stress of the machinery, not a reconstruction).

Usage (GPU box): python profiles/run_c5_densify.py > gpurun_out/c5_densify.json
"""
import json
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
from torch import nn  # noqa: E402

from gaussian_splatting_b200 import densify, rasterizer, synthetic  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    import diff_gaussian_rasterization as dgr
    cfg = synthetic.CONFIGS["C3"]
    gs, cam = synthetic.make_scene(cfg, P=500_000)
    H, W = cfg.H, cfg.W
    rs = dgr.GaussianRasterizationSettings(
        image_height=H, image_width=W, tanfovx=float(cam["tanfovx"]), tanfovy=float(cam["tanfovy"]),
        bg=torch.zeros(3, device=dev), scale_modifier=1.0, viewmatrix=cam["viewmatrix"].to(dev),
        projmatrix=cam["projmatrix"].to(dev), sh_degree=3, campos=cam["campos"].to(dev), prefiltered=False, debug=False,
        antialiasing=True)
    rast = dgr.GaussianRasterizer(rs)
    gm = types.SimpleNamespace(percent_dense=0.01)
    raw = {"xyz": gs["means3D"], "f_dc": gs["shs"][:, :1].contiguous(), "f_rest": gs["shs"][:, 1:].contiguous(),
           "opacity": torch.logit(gs["opacities"].clamp(1e-4, 1 - 1e-4)), "scaling": torch.log(gs["scales"]),
           "rotation": gs["rotations"]}
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 1.25e-4, "opacity": 0.025, "scaling": 0.005, "rotation": 0.001}
    groups = []
    for name in densify.GROUPS:
        p = nn.Parameter(raw[name].to(dev))
        setattr(gm, densify._ATTR[name], p)
        groups.append({"params": [p], "lr": lrs[name], "name": name})
    gm.optimizer = dgr.SparseGaussianAdam(groups, lr=0.0, eps=1e-15)
    P = gm._xyz.shape[0]
    gm.xyz_gradient_accum, gm.denom = torch.zeros(P, 1, device=dev), torch.zeros(P, 1, device=dev)
    gm.max_radii2D = torch.zeros(P, device=dev)
    target, _ = synthetic.make_targets(W, H, 3)
    target = target.to(dev)
    gen = torch.Generator(device=dev).manual_seed(0)

    def step():
        m2 = torch.zeros_like(gm._xyz, requires_grad=True)
        color, radii, _ = rast(means3D=gm._xyz, means2D=m2, opacities=torch.sigmoid(gm._opacity), dc=gm._features_dc,
                               shs=gm._features_rest, scales=torch.exp(gm._scaling),
                               rotations=torch.nn.functional.normalize(gm._rotation))
        (color - target).abs().mean().backward()
        densify.add_densification_stats(gm, m2, None, radii=radii)
        gm.optimizer.step(radii > 0, radii.shape[0])
        gm.optimizer.zero_grad(set_to_none=True)
        return radii

    def ms_of(fn, reps=3):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            out = fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps, out

    rows = []
    extent = float(torch.exp(gm._scaling.detach()).max(1).values.median()) / 0.01     # half clone, half split
    while gm._xyz.shape[0] < 5_000_000 and len(rows) < 40:
        P0 = gm._xyz.shape[0]
        step()                                                       # warm-up at this size (workspace regrowth)
        ms_step, radii = ms_of(step)
        g = torch.nan_to_num(gm.xyz_gradient_accum / gm.denom, nan=0.0)[:, 0]
        thr = float(torch.quantile(g[g > 0][:1_000_000], 0.75))
        ms_d, info = ms_of(lambda: densify.densify_and_prune(gm, thr, 0.005, extent, 20, radii, generator=gen), reps=1)
        st = rasterizer.workspace_stats(dev)
        rows.append({"P": P0, "P_after": info["P"], "kept": info["kept"], "clones": info["clones"], "split": info["split"],
                     "train_step_ms": round(ms_step, 3), "densify_and_prune_ms": round(ms_d, 3),
                     "R": st["last_num_rendered"], "scratch_MB": round(st["scratch_bytes"] / 1e6, 1), "regrows": st["regrows"],
                     "allocated_MB": round(torch.cuda.memory_allocated(dev) / 1e6, 1)})
    print(json.dumps({"config": "C5 (1080p, growth by densify_and_prune)", "rounds": rows}))


if __name__ == "__main__":
    main()
