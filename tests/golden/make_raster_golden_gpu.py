#!/usr/bin/env python
"""Generates tests/golden/raster_golden.npz from the REFERENCE'S OWN CUDA rasterizer (oracle/_ref, built from
/root/reference by oracle/build_ref.py) — needs a GPU, so it runs on the B200 box:

    gpurun -- 'python tests/golden/make_raster_golden_gpu.py gpurun_out/raster_golden.npz'
    cp gpurun_out/raster_golden.npz tests/golden/            # then commit

For a few small seeded scenes (tests/helpers.small_scene, so the inputs are regenerated bit-identically from the seed
and need not be stored) it records what the reference computes: radii, num_rendered, sorted keys, point_list, tile
ranges, n_contrib, final_T, colour, inverse depth, and the gradients for seeded upstream gradients (mean of 3 runs: the
reference's atomics are order-dependent).  tests/test_raster_golden_cpu.py then pins the C oracle and the emulated
product kernels to the reference itself in the GPU-less container."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ref_loader  # noqa: E402
from tests.helpers import ref_forward_state, run_module, small_scene  # noqa: E402

# name -> small_scene keyword arguments (tests/test_raster_golden_cpu.py rebuilds the same scenes)
CASES = {
    "a": dict(P=600, W=64, H=48, seed=1, antialiasing=True, bg=(0.1, 0.3, 0.6)),
    "b": dict(P=300, W=48, H=32, seed=2, sh_degree=2),
    "c": dict(P=1500, W=37, H=21, seed=4, scale_mult=10.0, behind_fraction=0.0),
    "d": dict(P=33, W=17, H=5, seed=3, scale_mult=2.0, sh_degree=0, bg=(1.0, 1.0, 1.0)),
}


def upstream(W, H, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(3, H, W, generator=g) / 100, torch.randn(1, H, W, generator=g) / 100


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tests", "golden", "raster_golden.npz")
    out = {}
    for variant in ("restored", "asis"):
        ref = ref_loader.load(variant)
        for name, kw in CASES.items():
            gs, s = small_scene(**kw)
            dc, dd = upstream(kw["W"], kw["H"])
            if variant == "restored":
                st = ref_forward_state(ref, gs, s)
                for k in ("num_rendered", "color", "radii", "invdepth", "point_list", "point_list_keys", "ranges",
                          "n_contrib", "final_T"):
                    out[f"{name}_{k}"] = np.asarray(st[k])
            runs = [run_module(ref, gs, s, dc, dd)["grads"] for _ in range(3)]
            for k in runs[0]:
                out[f"{name}_{variant}_grad_{k}"] = np.mean([r[k] for r in runs], axis=0).astype(np.float32)
    np.savez_compressed(out_path, **out)
    print("wrote", out_path, len(out), "arrays")


if __name__ == "__main__":
    main()
