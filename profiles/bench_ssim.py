#!/usr/bin/env python
"""Fused SSIM on one B200: ours, the reference's ssim.cu (oracle/_ref/fused_ssim) and the
conv2d formulation of utils/loss_utils.py, at the training image sizes and at the fused-ssim test's size
(tests/test.py:59: B, CH, H, W = 5, 5, 1080, 1920).  CUDA events, 5 warm-up + 30 timed calls; between calls
the working set is touched by the backward so nothing but L2 residency carries over (sizes >= 75 MB each way).

Algorithmic HBM bytes per pixel-channel: forward 8 read (img1, img2) + 12 written (three derivative maps),
backward 20 read (three maps, img1, img2) + 4 written = 44 B for the fused-mean path (the map variant adds
4 B written + 4 B read).  Usage (GPU box): python profiles/bench_ssim.py > gpurun_out/ssim.json
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402


def timeit(fn, warmup=5, reps=30):
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


def conv_ssim(window):
    def f(img1, img2):
        ch = img1.size(-3)
        w = window.expand(ch, 1, 11, 11).contiguous()
        mu1, mu2 = F.conv2d(img1, w, padding=5, groups=ch), F.conv2d(img2, w, padding=5, groups=ch)
        s1 = F.conv2d(img1 * img1, w, padding=5, groups=ch) - mu1.pow(2)
        s2 = F.conv2d(img2 * img2, w, padding=5, groups=ch) - mu2.pow(2)
        s12 = F.conv2d(img1 * img2, w, padding=5, groups=ch) - mu1 * mu2
        return (((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1.pow(2) + mu2.pow(2) + 1e-4) * (s1 + s2 + 9e-4))).mean()
    return f


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    import fused_ssim as ours
    impls = {"ours": ours.fused_ssim}
    only_ours = "--ours-only" in sys.argv          # A/B runs of kernel variants (GSR_LIB=...)
    try:
        if only_ours:
            raise RuntimeError("skipped (--ours-only)")
        from oracle import ref_loader
        impls["reference_cuda"] = ref_loader.load_fused_ssim().fused_ssim
    except Exception as ex:  # noqa: BLE001
        print(f"# reference fused-ssim unavailable: {ex}", file=sys.stderr)
    from oracle import ssim_ref
    g = torch.from_numpy(ssim_ref.taps()).to(dev)
    if not only_ours:
        impls["conv2d"] = conv_ssim((g[:, None] @ g[None, :])[None, None])
    peak = 6552.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:  # noqa: BLE001
        pass
    out = {"hbm_peak_GBps": peak, "shapes": {}}
    for shape in ((1, 3, 1080, 1920), (1, 3, 840, 1297), (1, 3, 2160, 3840), (5, 5, 1080, 1920)):
        gen = torch.Generator().manual_seed(0)
        img1 = torch.rand(shape, generator=gen).to(dev).requires_grad_(True)
        img2 = torch.rand(shape, generator=gen).to(dev)
        n = img1.numel()
        row = {}
        for name, fn in impls.items():
            def fwd():
                with torch.no_grad():
                    fn(img1, img2)

            def fwdbwd():
                img1.grad = None
                fn(img1, img2).backward()
            r = {"fwd_ms": round(timeit(fwd), 4), "fwd_bwd_ms": round(timeit(fwdbwd), 4)}
            if name == "ours":
                def infer():
                    with torch.no_grad():
                        fn(img1, img2, train=False)
                r["inference_ms"] = round(timeit(infer), 4)
                r["fwd_bwd_GBps"] = round(44 * n / (r["fwd_bwd_ms"] * 1e-3) / 1e9, 1)
                r["fwd_bwd_frac_of_hbm_peak"] = round(r["fwd_bwd_GBps"] / peak, 3)
            row[name] = r
        if "reference_cuda" in row:
            row["speedup_vs_reference_cuda"] = round(row["reference_cuda"]["fwd_bwd_ms"] / row["ours"]["fwd_bwd_ms"], 2)
        if "conv2d" in row:
            row["speedup_vs_conv2d"] = round(row["conv2d"]["fwd_bwd_ms"] / row["ours"]["fwd_bwd_ms"], 2)
        out["shapes"]["x".join(map(str, shape))] = row
        print(f"# {shape}: {row}", file=sys.stderr)
        del img1, img2
        torch.cuda.empty_cache()
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
