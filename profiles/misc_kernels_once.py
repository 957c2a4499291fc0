#!/usr/bin/env python
"""The kernels outside the rasterizer's hot path, each run a few times at benchmark size (the ncu target for
csrc/ssim.cu incl. the fused training loss, csrc/adam.cu, csrc/densify.cu, the PLY row kernels, and the two halves of the
view-parallel exchange — sender branch of preprocess_bwd_kernel + exchange_expand_kernel — on ONE GPU with the
cross-rank sum left out: ncu replays kernels, so it cannot run under a multi-rank launch with rank barriers; NVLink
byte counters therefore do not exist for these two).  Usage: python profiles/misc_kernels_once.py [reps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import diff_gaussian_rasterization as dgr  # noqa: E402
import fused_ssim as fs  # noqa: E402
from gaussian_splatting_b200 import loss as L, rasterizer, synthetic  # noqa: E402
from tests.test_exchange_gpu import LocalExchange  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
dev = torch.device("cuda", 0)
g = torch.Generator().manual_seed(0)
H, W = 1080, 1920
img = torch.rand(3, H, W, generator=g).to(dev).requires_grad_(True)
gt = torch.rand(3, H, W, generator=g).to(dev)
invd = torch.rand(1, H, W, generator=g).to(dev).requires_grad_(True)
mono = torch.rand(1, H, W, generator=g).to(dev)
for _ in range(reps):
    img.grad = None
    fs.fused_ssim(img[None], gt[None]).backward()                                  # ssim_fwd / ssim_mean / ssim_bwd
    img.grad = invd.grad = None
    L.training_loss(img, gt, 0.2, invdepth=invd, mono_invdepth=mono, depth_weight=0.1).backward()   # loss_fwd / finish / bwd

# sparse Adam over the six tensors of a 3M-Gaussian model
P = 3_000_000
params = [torch.nn.Parameter(torch.randn(P, w, device=dev)) for w in (3, 3, 45, 1, 3, 4)]
opt = dgr.SparseGaussianAdam([{"params": [p], "lr": 1e-3} for p in params], lr=0.0, eps=1e-15)
vis = torch.rand(P, device=dev) > 0.3
for _ in range(reps):
    for p in params:
        p.grad = torch.randn_like(p)
    opt.step(vis, P)

# the exchange at C3 on one GPU (plain stores)
cfg = synthetic.CONFIGS["C3"]
gs, cam = synthetic.make_scene(cfg)
s = dict(cam)
s.update(bg=torch.zeros(3), scale_modifier=1.0, sh_degree=3, prefiltered=False, debug=False)
from tests.helpers import settings_for  # noqa: E402
leaves = {k: v.to(dev).requires_grad_(True) for k, v in gs.items()}
m2 = torch.zeros(cfg.P, 3, device=dev, requires_grad=True)
rast = dgr.GaussianRasterizer(settings_for(dgr, s, dev))
dc = torch.rand(3, cfg.H, cfg.W, device=dev) / (cfg.W * cfg.H)
dd = torch.rand(1, cfg.H, cfg.W, device=dev) / (cfg.W * cfg.H)
rasterizer.set_gradient_reducer(LocalExchange())
try:
    for _ in range(reps):
        for t in list(leaves.values()) + [m2]:
            t.grad = None
        color, radii, invdepth = rast(means3D=leaves["means3D"], means2D=m2, opacities=leaves["opacities"], shs=leaves["shs"],
                                      scales=leaves["scales"], rotations=leaves["rotations"])
        torch.autograd.backward([color, invdepth], [dc, dd])
finally:
    rasterizer.set_gradient_reducer(None)
torch.cuda.synchronize()
print("ok")
