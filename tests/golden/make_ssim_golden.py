#!/usr/bin/env python
"""Generates tests/golden/ssim_golden.npz from the REFERENCE's own Python SSIM.

Run in the authoring container only (needs /root/reference; the GPU box and CI never run this):
    python tests/golden/make_ssim_golden.py
It imports /root/reference/utils/loss_utils.py unmodified and evaluates `ssim(img1, img2)`
(loss_utils.py:56-86: F.conv2d with the 11x11 Gaussian window, zero padding — the formulation the
fused-ssim submodule's tests/test.py:57-91 asserts its CUDA kernel against) on CPU in float64 and
float32, with the gradient w.r.t. img1 from torch.autograd.  The "valid" variant is the same map
cropped by 5 pixels before the mean (fused_ssim/__init__.py:13-14), computed here from the
reference's `_ssim` building blocks.  Also stores the window of `create_window`.
"""
import importlib.util
import os

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("ref_loss_utils", "/root/reference/utils/loss_utils.py")
lu = importlib.util.module_from_spec(spec)
spec.loader.exec_module(lu)


def ssim_map(img1, img2):
    """The body of loss_utils._ssim (:68-81) without the final mean, to crop for padding='valid'."""
    ch = img1.size(-3)
    w = lu.create_window(11, ch).type_as(img1)
    mu1, mu2 = F.conv2d(img1, w, padding=5, groups=ch), F.conv2d(img2, w, padding=5, groups=ch)
    s1 = F.conv2d(img1 * img1, w, padding=5, groups=ch) - mu1.pow(2)
    s2 = F.conv2d(img2 * img2, w, padding=5, groups=ch) - mu2.pow(2)
    s12 = F.conv2d(img1 * img2, w, padding=5, groups=ch) - mu1 * mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1.pow(2) + mu2.pow(2) + C1) * (s1 + s2 + C2))


def main():
    g = torch.Generator().manual_seed(20)
    out = {"window_1d": lu.gaussian(11, 1.5).numpy(), "window_2d": lu.create_window(11, 1)[0, 0].numpy()}
    cases = {"a": (2, 3, 37, 53), "b": (1, 1, 11, 11), "c": (1, 3, 33, 70), "d": (1, 2, 12, 75)}
    for name, shape in cases.items():
        img1 = torch.rand(shape, generator=g, dtype=torch.float64)
        # correlated target: img2 = noisy, blurred img1 (SSIM away from 0), as a render vs. its ground truth
        img2 = (0.7 * img1 + 0.3 * torch.rand(shape, generator=g, dtype=torch.float64)).clamp(0, 1)
        out[f"{name}_img1"], out[f"{name}_img2"] = img1.float().numpy(), img2.float().numpy()
        for tag, dt in (("f64", torch.float64), ("f32", torch.float32)):
            x = img1.float().to(dt).requires_grad_(True)     # identical fp32-representable inputs in both precisions
            y = img2.float().to(dt)
            v = lu.ssim(x, y)                                 # the reference's own entry point
            v.backward()
            out[f"{name}_same_value_{tag}"] = v.detach().numpy()
            if tag == "f64":
                out[f"{name}_same_grad_{tag}"] = x.grad.numpy()
            x2 = img1.float().to(dt).requires_grad_(True)
            m = ssim_map(x2, y)
            assert torch.allclose(m.mean(), v.detach(), rtol=1e-6 if dt == torch.float32 else 1e-12)
            if tag == "f64":
                out[f"{name}_map_f64"] = m.detach().numpy()
            if shape[2] > 10 and shape[3] > 10:
                vv = m[:, :, 5:-5, 5:-5].mean()
                vv.backward()
                out[f"{name}_valid_value_{tag}"] = vv.detach().numpy()
                if tag == "f64":
                    out[f"{name}_valid_grad_{tag}"] = x2.grad.numpy()
    np.savez_compressed(os.path.join(HERE, "ssim_golden.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
