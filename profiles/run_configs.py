#!/usr/bin/env python
"""All BASELINE.json configs on one B200, ours next to the reference's CUDA rasterizer (oracle/_ref):

  C1  100K Gaussians, 800x800: fwd+bwd
  C2  1M Gaussians, 1297x840: FULL TRAIN STEP = activations (exp / sigmoid / normalize / cat as
      scene/gaussian_model.py:130-157), rasterize, L1 losses, backward, torch.optim.Adam(eps=1e-15)
      with the six parameter groups of gaussian_model.py:235-245
  C3  3M Gaussians, 1920x1080, antialiasing: forward-only FPS (and fwd+bwd, the bench.py headline)
  C4  6M Gaussians, 3840x2160: one camera on one GPU (the 8-camera batch is bench.py --gpus 8 --workload C4)
  C5  densification stress: P = 500K * 10^(k/10), k = 0..10 at 1080p; time, R, scratch regrowth

Usage (GPU box): python profiles/run_configs.py > gpurun_out/configs.json
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from gaussian_splatting_b200 import rasterizer, synthetic  # noqa: E402


def settings(module, s, dev):
    return module.GaussianRasterizationSettings(
        image_height=int(s["image_height"]), image_width=int(s["image_width"]), tanfovx=float(s["tanfovx"]),
        tanfovy=float(s["tanfovy"]), bg=torch.zeros(3, device=dev), scale_modifier=1.0,
        viewmatrix=s["viewmatrix"].to(dev), projmatrix=s["projmatrix"].to(dev), sh_degree=3,
        campos=s["campos"].to(dev), prefiltered=False, debug=False, antialiasing=bool(s["antialiasing"]))


def timeit(fn, warmup=3, reps=10):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def scene(cfg, P=None):
    gs, cam = synthetic.make_scene(cfg, P=P or cfg.P)
    return gs, dict(cam)


def fwd_bwd_fns(module, gs, s, dev):
    params = {k: v.to(dev).requires_grad_(True) for k, v in gs.items()}
    P = gs["means3D"].shape[0]
    m2 = torch.zeros(P, 3, device=dev, requires_grad=True)
    H, W = int(s["image_height"]), int(s["image_width"])
    g = torch.Generator().manual_seed(5)
    dc = ((torch.rand(3, H, W, generator=g) - 0.5) / (3 * H * W)).to(dev)
    dd = ((torch.rand(1, H, W, generator=g) - 0.5) * 0.1 / (H * W)).to(dev)
    rast = module.GaussianRasterizer(settings(module, s, dev))

    def fwd():
        with torch.no_grad():
            rast(means3D=params["means3D"], means2D=m2, opacities=params["opacities"], shs=params["shs"],
                 scales=params["scales"], rotations=params["rotations"])

    def fwdbwd():
        for t in list(params.values()) + [m2]:
            t.grad = None
        c, r, d = rast(means3D=params["means3D"], means2D=m2, opacities=params["opacities"], shs=params["shs"],
                       scales=params["scales"], rotations=params["rotations"])
        torch.autograd.backward([c, d], [dc, dd])

    return fwd, fwdbwd


def train_step_fn(module, gs, s, dev, sparse_adam=False, ssim=None, fused_act=False, fused_loss=False):
    """C2: the per-iteration work of train.py:143-251 without datasets.
    sparse_adam: the fork's accelerated branch (separate_sh / dc= keyword, optimizer.step(visible, N)).
    ssim: a fused_ssim(img1, img2) callable -> the full loss of train.py:155-162,
    0.8 * L1 + 0.2 * (1 - SSIM), plus the inverse-depth L1 term.
    fused_act: the kernels apply exp / normalize / sigmoid themselves (raw_inputs, ours only).
    fused_loss: gaussian_splatting_b200.loss.training_loss — the whole loss in one forward and one backward launch."""
    P = gs["means3D"].shape[0]
    nnp = torch.nn.Parameter
    xyz = nnp(gs["means3D"].to(dev))
    f_dc = nnp(gs["shs"][:, :1, :].contiguous().to(dev))
    f_rest = nnp(gs["shs"][:, 1:, :].contiguous().to(dev))
    scaling = nnp(torch.log(gs["scales"]).to(dev))
    rotation = nnp(gs["rotations"].to(dev))
    opacity = nnp(torch.logit(gs["opacities"].clamp(1e-4, 1 - 1e-4)).to(dev))
    Opt = module.SparseGaussianAdam if sparse_adam else torch.optim.Adam
    opt = Opt([
        {"params": [xyz], "lr": 1.6e-4}, {"params": [f_dc], "lr": 2.5e-3}, {"params": [f_rest], "lr": 2.5e-3 / 20},
        {"params": [opacity], "lr": 0.025}, {"params": [scaling], "lr": 0.005}, {"params": [rotation], "lr": 0.001},
    ], lr=0.0, eps=1e-15)
    H, W = int(s["image_height"]), int(s["image_width"])
    target, target_d = synthetic.make_targets(W, H, 2)
    target, target_d = target.to(dev), target_d.to(dev)
    rast = module.GaussianRasterizer(settings(module, s, dev))

    def step():
        m2 = torch.zeros(P, 3, device=dev, requires_grad=True)
        kw = dict(dc=f_dc, shs=f_rest) if sparse_adam else dict(shs=torch.cat((f_dc, f_rest), dim=1))
        if fused_act:
            color, radii, invd = rast(means3D=xyz, means2D=m2, opacities=opacity, scales=scaling, rotations=rotation,
                                      raw_inputs=7, **kw)
        else:
            color, radii, invd = rast(means3D=xyz, means2D=m2, opacities=torch.sigmoid(opacity),
                                      scales=torch.exp(scaling), rotations=torch.nn.functional.normalize(rotation), **kw)
        if fused_loss:
            from gaussian_splatting_b200 import loss as _loss
            loss = _loss.training_loss(color, target, 0.2, invdepth=invd, mono_invdepth=target_d, depth_weight=0.1)
        elif ssim is None:
            loss = (color - target).abs().mean() + 0.1 * (invd - target_d).abs().mean()
        else:
            loss = 0.8 * (color - target).abs().mean() + 0.2 * (1.0 - ssim(color.unsqueeze(0), target.unsqueeze(0))) \
                + 0.1 * (invd - target_d).abs().mean()
        loss.backward()
        if sparse_adam:
            opt.step(radii > 0, radii.shape[0])               # train.py:240-247
        else:
            opt.step()
        opt.zero_grad(set_to_none=True)

    return step


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    import diff_gaussian_rasterization as ours
    mods = {"ours": ours}
    try:
        from oracle import ref_loader
        mods["reference"] = ref_loader.load("restored")
    except Exception as ex:  # noqa: BLE001
        print(f"# reference CUDA build unavailable: {ex}", file=sys.stderr)
    ssim_of = {}
    try:
        import fused_ssim
        ssim_of["ours"] = fused_ssim.fused_ssim
        from oracle import ref_loader
        ssim_of["reference"] = ref_loader.load_fused_ssim().fused_ssim
    except Exception as ex:  # noqa: BLE001
        print(f"# fused-ssim unavailable: {ex}", file=sys.stderr)
    out = {}
    # C5: densification stress (ours): growing P, dynamic workspace, sort bit-width via resolution
    c3 = synthetic.CONFIGS["C3"]
    ours = mods["ours"]
    c5 = []
    for k in range(11):
        P = int(500_000 * 10 ** (k / 10))
        gs, s = scene(c3, P=P)
        fwd, fwdbwd = fwd_bwd_fns(ours, gs, s, dev)
        t0 = time.time()
        ms = timeit(fwdbwd, warmup=1, reps=3)
        st = rasterizer.workspace_stats(dev)
        c5.append({"P": P, "fwd_bwd_ms": round(ms, 3), "R": st["last_num_rendered"], "scratch_MB": round(st["scratch_bytes"] / 1e6, 1),
                   "regrows": st["regrows"], "wall_s": round(time.time() - t0, 2)})
        torch.cuda.empty_cache()
    out["C5"] = c5
    for name in ("C1", "C2", "C3", "C4"):
        cfg = synthetic.CONFIGS[name]
        gs, s = scene(cfg)
        row = {"P": cfg.P, "W": cfg.W, "H": cfg.H}
        for mname, m in mods.items():
            fwd, fwdbwd = fwd_bwd_fns(m, gs, s, dev)
            r = {"fwd_ms": round(timeit(fwd), 3), "fwd_bwd_ms": round(timeit(fwdbwd), 3)}
            if name == "C2":
                r["train_step_ms"] = round(timeit(train_step_fn(m, gs, s, dev)), 3)
                if hasattr(m, "SparseGaussianAdam"):
                    r["train_step_sparse_adam_ms"] = round(timeit(train_step_fn(m, gs, s, dev, sparse_adam=True)), 3)
                if mname in ssim_of:
                    # the full loss of train.py:155-162 with each side's fastest shipped pieces:
                    # ours = SparseGaussianAdam + our fused SSIM, reference = torch Adam + its own fused-ssim CUDA
                    r["train_step_ssim_ms"] = round(timeit(train_step_fn(
                        m, gs, s, dev, sparse_adam=hasattr(m, "SparseGaussianAdam"), ssim=ssim_of[mname])), 3)
                if mname == "ours":
                    r["train_step_sparse_adam_fused_act_ms"] = round(timeit(train_step_fn(
                        m, gs, s, dev, sparse_adam=True, fused_act=True)), 3)
                    r["train_step_ssim_fused_act_ms"] = round(timeit(train_step_fn(
                        m, gs, s, dev, sparse_adam=True, ssim=ssim_of.get("ours"), fused_act=True)), 3)
                    try:
                        r["train_step_fused_act_fused_loss_ms"] = round(timeit(train_step_fn(
                            m, gs, s, dev, sparse_adam=True, fused_act=True, fused_loss=True)), 3)
                    except Exception as ex:  # noqa: BLE001
                        r["train_step_fused_act_fused_loss_ms"] = f"unavailable: {ex}"
            row[mname] = r
            torch.cuda.empty_cache()
        if "reference" in row:
            row["speedup_fwd_bwd"] = round(row["reference"]["fwd_bwd_ms"] / row["ours"]["fwd_bwd_ms"], 2)
            row["speedup_fwd"] = round(row["reference"]["fwd_ms"] / row["ours"]["fwd_ms"], 2)
        out[name] = row
        print(f"# {name}: {row}", file=sys.stderr)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
