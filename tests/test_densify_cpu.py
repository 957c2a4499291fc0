"""CPU tests of the density-control oracle (oracle/densify_ref.py) against golden vectors generated from the
reference's own GaussianModel (scene/gaussian_model.py:323-327,400-633, train.py:212-215) by
tests/golden/make_densify_golden.py.  Structure (which rows survive, in which order) must be exact; values the
reference only copies must be bit-identical; values it recomputes (children's xyz and scaling, reset opacity) are
compared to fp32 rounding."""
import os

import numpy as np
import pytest

from oracle import densify_ref as dr

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "densify_golden.npz"))
CASES = ("a", "b", "c", "d", "e")
STATE_KEYS = [pre + g for g in dr.GROUPS for pre in ("", "m_", "v_")] + ["xyz_gradient_accum", "denom", "max_radii2D"]


def load_state(prefix):
    return {k: GOLD[prefix + k].copy() for k in STATE_KEYS}


def assert_state_equal(got, want_prefix, recomputed=()):
    for k in STATE_KEYS:
        w = GOLD[want_prefix + k]
        assert got[k].shape == w.shape, (k, got[k].shape, w.shape)
        assert got[k].dtype == w.dtype, (k, got[k].dtype, w.dtype)
        if k in recomputed:
            np.testing.assert_allclose(got[k], w, rtol=2e-6, atol=2e-6, err_msg=k)
        else:
            assert np.array_equal(got[k], w), k


def densify_case(case):
    s = load_state(f"{case}_in_")
    max_grad, min_opacity, extent, size_thr, pd = GOLD[f"{case}_args"]
    out, info = dr.densify_and_prune(s, max_grad, min_opacity, extent, size_thr if size_thr else None,
                                     GOLD[f"{case}_radii"], pd, GOLD[f"{case}_noise"])
    return s, out, info


@pytest.mark.parametrize("case", CASES)
def test_statistics_match_reference(case):
    P = GOLD[f"{case}_in_xyz"].shape[0]
    s = {"xyz_gradient_accum": np.zeros((P, 1), np.float32), "denom": np.zeros((P, 1), np.float32),
         "max_radii2D": np.zeros((P,), np.float32)}
    for v in range(4):
        dr.add_densification_stats(s, GOLD[f"{case}_view{v}_grad"], GOLD[f"{case}_view{v}_vis"],
                                   GOLD[f"{case}_view{v}_radii"])
    np.testing.assert_allclose(s["xyz_gradient_accum"], GOLD[f"{case}_in_xyz_gradient_accum"], rtol=3e-7, atol=0)
    assert np.array_equal(s["denom"], GOLD[f"{case}_in_denom"])
    assert np.array_equal(s["max_radii2D"], GOLD[f"{case}_in_max_radii2D"])


@pytest.mark.parametrize("case", CASES)
def test_densify_and_prune_matches_reference(case):
    s, out, info = densify_case(case)
    assert_state_equal(out, f"{case}_out_", recomputed=("xyz", "scaling"))
    # the statistics are reset by densification_postfix (gaussian_model.py:506-508)
    assert not out["xyz_gradient_accum"].any() and not out["denom"].any() and not out["max_radii2D"].any()
    assert info["n_clone"] + info["n_split"] <= s["xyz"].shape[0]


def test_cases_exercise_every_branch():
    counts = {c: densify_case(c)[2] for c in CASES}
    assert counts["a"]["n_clone"] > 0 and counts["a"]["n_split"] > 0 and counts["a"]["n_pruned"] > 0
    assert counts["b"]["n_clone"] > 0 and counts["b"]["n_split"] > 0
    assert counts["c"]["n_clone"] == 0 and counts["c"]["n_split"] == 0 and counts["c"]["n_pruned"] > 0
    assert counts["d"]["n_clone"] == 0 and counts["d"]["n_split"] > 0


def test_reset_opacity_and_prune_points_match_reference():
    s = load_state("a_out_")
    dr.reset_opacity(s)
    assert_state_equal(s, "a_reset_", recomputed=("opacity",))
    assert not s["m_opacity"].any() and not s["v_opacity"].any()
    s = load_state("a_reset_")
    s, _ = dr.prune_points(s, GOLD["a_prune_mask"])
    assert_state_equal(s, "a_pruned_")


def test_order_of_the_result():
    """[surviving originals][clones][first children][second children], each in source order — the layout the
    fused kernel must reproduce (cat at gaussian_model.py:470-484, prune at :549-550, :611)."""
    s, out, info = densify_case("a")
    clone, split = info["clone"], info["split"]
    staged = np.concatenate((s["opacity"][~split], s["opacity"][clone], s["opacity"][split], s["opacity"][split]))
    # the result is a subsequence of the staged layout (the final prune only removes rows)
    j = 0
    for row in out["opacity"][:, 0]:
        while staged[j, 0] != row:
            j += 1
        j += 1
    assert j <= staged.shape[0]
