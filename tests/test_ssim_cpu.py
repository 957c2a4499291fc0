"""CPU tests of the SSIM oracle (oracle/ssim_ref.py) against golden vectors generated from the
reference's own Python SSIM (utils/loss_utils.py:56-86) by tests/golden/make_ssim_golden.py —
the formulation the fused-ssim submodule's own test (tests/test.py:57-91) asserts its kernel against."""
import os

import numpy as np
import pytest

from oracle import ssim_ref

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ssim_golden.npz"))
CASES = ("a", "b", "c", "d")


def test_window_is_the_references():
    # create_window(11, 1) of loss_utils.py:46-54 == outer product of the taps ssim.cu:9-19 hard-codes
    g = ssim_ref.taps(np.float32)
    assert np.array_equal(g, GOLD["window_1d"])
    assert np.array_equal(np.outer(g, g).astype(np.float32), GOLD["window_2d"])
    assert abs(float(g.astype(np.float64).sum()) - 1.0) < 1e-7 and np.array_equal(g, g[::-1])


@pytest.mark.parametrize("case", CASES)
def test_map_matches_reference_fp64(case):
    m, *_ = ssim_ref.ssim_forward(GOLD[f"{case}_img1"], GOLD[f"{case}_img2"], dtype=np.float64)
    # the reference's conv2d uses the 2-D window rounded to fp32 (create_window: `.mm(...).float()`), the fused
    # kernel the unrounded product of the fp32 taps: the two reference formulations differ by ~2e-8 themselves
    np.testing.assert_allclose(m, GOLD[f"{case}_map_f64"], rtol=0, atol=1e-7)


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("padding", ("same", "valid"))
def test_value_and_gradient_match_reference_autograd(case, padding):
    img1, img2 = GOLD[f"{case}_img1"], GOLD[f"{case}_img2"]
    if padding == "valid" and (img1.shape[2] <= 10 or img1.shape[3] <= 10):
        pytest.skip("empty crop")
    v, g = ssim_ref.fused_ssim(img1, img2, padding, dtype=np.float64)
    assert abs(v - GOLD[f"{case}_{padding}_value_f64"]) < 1e-8
    # the analytic derivative maps (ssim.cu:268-277) + window sums == autograd through conv2d
    ref = GOLD[f"{case}_{padding}_grad_f64"]
    np.testing.assert_allclose(g, ref, rtol=0, atol=2e-7 * np.abs(ref).max())
    # fp32 restatement (the kernels' precision) stays within fp32 noise of the truth
    v32, g32 = ssim_ref.fused_ssim(img1, img2, padding, dtype=np.float32)
    assert abs(float(v32) - float(GOLD[f"{case}_{padding}_value_f32"])) < 2e-6
    assert np.abs(g32 - ref).max() < 2e-5 * np.abs(ref).max() + 1e-9


def test_edge_shapes():
    rng = np.random.default_rng(3)
    for shape in ((1, 1, 1, 1), (1, 2, 3, 40), (2, 1, 33, 2)):
        a, b = rng.random(shape, dtype=np.float32), rng.random(shape, dtype=np.float32)
        m, d1, d2, d3 = ssim_ref.ssim_forward(a, b, dtype=np.float64)
        assert m.shape == shape and np.isfinite(m).all() and (m <= 1 + 1e-12).all()
        # identical images: ssim == 1 everywhere, gradient of the mean vanishes
        m1, *_ = ssim_ref.ssim_forward(a, a, dtype=np.float64)
        np.testing.assert_allclose(m1, 1.0, atol=1e-12)
        _, g = ssim_ref.fused_ssim(a, a, dtype=np.float64)
        assert np.abs(g).max() < 1e-9
    v, g = ssim_ref.fused_ssim(np.zeros((1, 1, 8, 8), np.float32), np.zeros((1, 1, 8, 8), np.float32), "valid")
    assert np.isnan(v) and not g.any()
