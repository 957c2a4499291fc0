"""Multi-GPU check of the fused (in-kernel, NVLink multicast) gradient reduction against a plain
NCCL all-reduce of independently computed gradients.  Needs >= 2 GPUs:

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/multi_gpu_check.py

Every rank renders its own camera of a ring over the same Gaussians; prints max relative errors.
"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    import diff_gaussian_rasterization as dgr
    from gaussian_splatting_b200 import synthetic
    from gaussian_splatting_b200.view_parallel import FusedGradientReducer
    from tests.helpers import settings_for

    P, W, H = 200_000, 640, 480
    gs = synthetic.make_gaussians(P, W, H, seed=4, sh_degree=3, scale_mult=1.5)
    cam = synthetic.make_camera(W, H, 0.9, 360.0 / world * rank)
    s = dict(cam)
    s.update(bg=torch.tensor([0.1, 0.2, 0.3]), scale_modifier=1.0, sh_degree=3, antialiasing=True,
             prefiltered=False, debug=False)
    g = torch.Generator().manual_seed(11)
    dc = torch.randn(3, H, W, generator=g).to(dev) / (W * H) ** 0.5
    dd = torch.randn(1, H, W, generator=g).to(dev) / (W * H) ** 0.5

    def run():
        leaves = {k: v.to(dev).requires_grad_(True) for k, v in gs.items()}
        m2 = torch.zeros(P, 3, device=dev, requires_grad=True)
        color, radii, invd = dgr.GaussianRasterizer(settings_for(dgr, s, dev))(
            means3D=leaves["means3D"], means2D=m2, opacities=leaves["opacities"], shs=leaves["shs"],
            scales=leaves["scales"], rotations=leaves["rotations"])
        torch.autograd.backward([color, invd], [dc, dd])
        return {k: v.grad.clone() for k, v in leaves.items()}, m2.grad.clone()

    # reference: local gradients + NCCL all-reduce
    want, m2_local = run()
    for v in want.values():
        dist.all_reduce(v)
    # fused: reduced inside the kernel
    red = FusedGradientReducer().install()
    worst = 0.0
    # factored exchange (default) and full-gradient reduction, each with both orderings of the buffer clear +
    # rank barrier: overlapped with the forward (default) and at backward start
    for rep in range(8):
        red.factored = rep < 4
        red.early = rep % 4 < 2
        got, m2 = run()
        torch.cuda.synchronize()
        for k in want:
            err = float((got[k] - want[k]).abs().max() / want[k].abs().max().clamp_min(1e-20))
            worst = max(worst, err)
            assert err < 1e-5, (k, err)
        # per-view statistic: never reduced (equal to the single-rank value up to atomic ordering)
        err2 = float((m2 - m2_local).abs().max() / m2_local.abs().max().clamp_min(1e-20))
        assert err2 < 1e-5, ("dL_dmeans2D must stay local", err2)
    red.uninstall()
    print(f"rank {rank}/{world}: fused multimem reduction == NCCL all-reduce, max rel err {worst:.2e}, "
          f"{red.bytes_reduced / 1e6:.1f} MB reduced per backward", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
