#!/usr/bin/env python
"""Kernel-level breakdown of the C2 full train step (profiles/run_configs.py: train_step_fn with
SparseGaussianAdam + fused SSIM) with torch.profiler: which kernels fill the time between the rasterizer
calls.  Usage (GPU box): python profiles/train_step_trace.py > gpurun_out/train_trace.txt"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "profiles"))

import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

import run_configs as rc  # noqa: E402
from gaussian_splatting_b200 import synthetic  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    import diff_gaussian_rasterization as ours
    import fused_ssim
    cfg = synthetic.CONFIGS[sys.argv[1] if len(sys.argv) > 1 else "C2"]
    gs, s = rc.scene(cfg)
    step = rc.train_step_fn(ours, gs, s, dev, sparse_adam=True, ssim=fused_ssim.fused_ssim)
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    ms = rc.timeit(step, warmup=2, reps=10)
    n = 10
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for _ in range(n):
            step()
        torch.cuda.synchronize()
    rows = [(e.key, e.device_time_total / n, e.count / n) for e in prof.key_averages() if e.device_time_total > 0]
    rows.sort(key=lambda r: -r[1])
    tot = sum(r[1] for r in rows)
    print(f"step {ms:.3f} ms (CUDA events); kernel time sum {tot / 1e3:.3f} ms/step; {sum(r[2] for r in rows):.0f} launches/step")
    for k, us, c in rows[:45]:
        print(f"{us:9.1f} us  x{c:4.1f}  {k[:110]}")


if __name__ == "__main__":
    main()
