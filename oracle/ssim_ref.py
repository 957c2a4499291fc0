"""CPU restatement of the reference's fused-ssim kernels (submodules/fused-ssim/ssim.cu) in numpy.

TEST INFRASTRUCTURE ONLY: imported by tests/ and __graft_entry__.smoke(); the product path
(gaussian_splatting_b200.ssim over libgsrast.so) never touches this file.

Pinned (tests/test_ssim_cpu.py) against tests/golden/ssim_golden.npz, which
tests/golden/make_ssim_golden.py generated in the authoring container by importing the reference's
own Python SSIM (utils/loss_utils.py:42-86 — the conv2d formulation that fused-ssim's tests/test.py:57-91
uses as its truth) and differentiating it with autograd; on the GPU the CUDA path is additionally
compared with the reference's ssim.cu compiled for sm_100 (oracle/_ref/fused_ssim, oracle/build_ref.py).

Every function takes `dtype`: np.float32 follows the kernels' precision step by step (products
rounded before the window sums, as multiply_shared_mem / do_sq do, ssim.cu:62-81), np.float64 is
the truth used for tolerances.
"""
import numpy as np

# gaussian(11, 1.5) / sum in float32: what create_window (loss_utils.py:42-54) computes with torch and what
# ssim.cu:9-19 hard-codes as G_00..G_10 (tests/test_ssim_cpu.py checks both statements)
_TAPS = (0.001028380123898387, 0.0075987582094967365, 0.036000773310661316, 0.10936068743467331,
         0.21300552785396576, 0.26601171493530273, 0.21300552785396576, 0.10936068743467331,
         0.036000773310661316, 0.0075987582094967365, 0.001028380123898387)


def taps(dtype=np.float32):
    return np.asarray(_TAPS, dtype=np.float32).astype(dtype)


def window_sum(img, dtype=np.float32):
    """Separable 11x11 zero-padded correlation of [..., H, W]: horizontal pass then vertical pass,
    taps accumulated left to right (do_separable_conv_x / _y, ssim.cu:98-185)."""
    g = taps(dtype)
    a = np.asarray(img, dtype=dtype)
    H, W = a.shape[-2:]
    pad = [(0, 0)] * (a.ndim - 2)
    ax = np.pad(a, pad + [(0, 0), (5, 5)])
    hx = np.zeros_like(a)
    for k in range(11):
        hx = hx + g[k] * ax[..., :, k:k + W]
    ay = np.pad(hx, pad + [(5, 5), (0, 0)])
    out = np.zeros_like(a)
    for k in range(11):
        out = out + g[k] * ay[..., k:k + H, :]
    return out


def ssim_forward(img1, img2, C1=0.01 ** 2, C2=0.03 ** 2, dtype=np.float32):
    """fusedssimCUDA (ssim.cu:187-281) -> (ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12)."""
    a = np.asarray(img1, dtype=dtype)
    b = np.asarray(img2, dtype=dtype)
    C1, C2 = dtype(C1), dtype(C2)
    mu1, mu2 = window_sum(a, dtype), window_sum(b, dtype)
    sigma1_sq = window_sum(a * a, dtype) - mu1 * mu1
    sigma2_sq = window_sum(b * b, dtype) - mu2 * mu2
    sigma12 = window_sum(a * b, dtype) - mu1 * mu2
    mu1_sq, mu2_sq, mu1_mu2 = mu1 * mu1, mu2 * mu2, mu1 * mu2
    two = dtype(2)
    C = two * mu1_mu2 + C1                      # ssim.cu:255-262
    D = two * sigma12 + C2
    A = mu1_sq + mu2_sq + C1
    B = sigma1_sq + sigma2_sq + C2
    m = (C * D) / (A * B)
    dm_dmu1 = ((mu2 * two * D) / (A * B) - (mu2 * two * C) / (A * B)      # ssim.cu:268-277
               - (mu1 * two * C * D) / (A * A * B) + (mu1 * two * C * D) / (A * B * B))
    dm_dsigma1_sq = (-C * D) / (A * B * B)
    dm_dsigma12 = (two * C) / (A * B)
    return m, dm_dmu1, dm_dsigma1_sq, dm_dsigma12


def ssim_backward(img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, dtype=np.float32):
    """fusedssim_backwardCUDA (ssim.cu:283-365) -> dL_dimg1."""
    a = np.asarray(img1, dtype=dtype)
    b = np.asarray(img2, dtype=dtype)
    dl = np.asarray(dL_dmap, dtype=dtype)
    d = window_sum(np.asarray(dm_dmu1, dtype=dtype) * dl, dtype)
    d = d + a * dtype(2) * window_sum(np.asarray(dm_dsigma1_sq, dtype=dtype) * dl, dtype)
    d = d + b * window_sum(np.asarray(dm_dsigma12, dtype=dtype) * dl, dtype)
    return d


def fused_ssim(img1, img2, padding="same", dtype=np.float32):
    """fused_ssim() of fused_ssim/__init__.py:35-41 plus its autograd backward for an upstream
    gradient of 1: -> (mean of the map, d mean / d img1)."""
    m, d_mu1, d_s1, d_s12 = ssim_forward(img1, img2, dtype=dtype)
    dl = np.zeros_like(m)
    if padding == "valid":                       # __init__.py:13-14, 29-31
        crop = m[..., 5:-5, 5:-5]
        dl[..., 5:-5, 5:-5] = dtype(1) / dtype(max(crop.size, 1))
    else:
        crop = m
        dl[...] = dtype(1) / dtype(max(crop.size, 1))
    mean = crop.mean(dtype=np.float64) if crop.size else np.float64("nan")
    return dtype(mean), ssim_backward(img1, img2, dl, d_mu1, d_s1, d_s12, dtype)
