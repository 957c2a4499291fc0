"""GPU parity tests (run on the B200 box with -m gpu): libgsrast.so through the reference-shaped
Python API against (a) the CPU oracle and (b) the reference's own CUDA rasterizer built for
sm_100 (oracle/_ref, see oracle/build_ref.py), on the same seeded inputs.

Tolerances (BASELINE.json north_star): bit-exact sorted values (point_list) / tile ranges / radii /
num_rendered / n_contrib (and the reference's sorted keys as REBUILT from ranges, point_list and the records'
depth by a test hook: the product never materialises tile<<32|depth); image and inverse depth <= 1e-5 max-abs;
gradients <= 1e-3 norm-wise (||a-b||inf / ||b||inf) AND <= 1e-2 per element for every element above
1e-4 * ||b||inf (SURVEY.md 8c asks for the per-element check; 1e-4 is what the reference's own atomic-order
noise allows against the mean of 3 of its runs).
"""
import numpy as np
import pytest
import torch

from tests.helpers import (elem_rel_err, ours_forward_state, ref_forward_state, rel_err, run_module,
                           run_oracle, small_scene)

pytestmark = pytest.mark.gpu

IMG_TOL = 1e-5
GRAD_TOL = 1e-3


@pytest.fixture(scope="module")
def dgr():
    if not torch.cuda.is_available():
        pytest.fail("CUDA device required for -m gpu tests")
    import diff_gaussian_rasterization as m
    from gaussian_splatting_b200 import _lib
    _lib.load()                       # fails loudly if libgsrast.so is missing
    return m


@pytest.fixture(scope="module")
def ref_restored():
    from oracle import ref_loader
    return ref_loader.load("restored")


@pytest.fixture(scope="module")
def ref_asis():
    from oracle import ref_loader
    return ref_loader.load("asis")


def _upstream(W, H, seed=5, depth=True):
    g = torch.Generator().manual_seed(seed)
    dc = torch.randn(3, H, W, generator=g) / (W * H) ** 0.5
    dd = torch.randn(1, H, W, generator=g) / (W * H) ** 0.5 if depth else None
    return dc, dd


ELEM_FLOOR = 1e-4


def _check_grads(got, want, keys, tol=GRAD_TOL, elem_floor=ELEM_FLOOR, outliers=0.0):
    """Norm-wise bound on every tensor, and per element (above elem_floor * ||b||inf) 10 * tol.  outliers: fraction of
    the checked elements that may miss the per-element bound (by less than 10x).  Only the full-size tests use it
    (1e-6 of up to 3.5e8 elements): there the reference's own atomic-order noise, of which the mean of a few runs
    still carries a part, reaches 1e-2 on an element near the floor about once in several runs."""
    for k in keys:
        a, b = np.asarray(got[k], np.float64), np.asarray(want[k], np.float64).reshape(got[k].shape)
        assert rel_err(a, b) < tol, (k, rel_err(a, b))
        if outliers == 0.0:
            assert elem_rel_err(a, b, elem_floor) < 10 * tol, (k, elem_rel_err(a, b, elem_floor))
            continue
        m = np.abs(b) > elem_floor * np.abs(b).max()
        rel = np.abs(a - b)[m] / np.abs(b)[m]
        bad = int((rel >= 10 * tol).sum())
        assert bad <= outliers * rel.size and (rel.max() if rel.size else 0.0) < 100 * tol, (k, bad, rel.size, float(rel.max()))


# ---------------------------------------------------------------- vs the CPU oracle
@pytest.mark.parametrize("W,H", [(48, 40), (67, 53), (160, 96)])
@pytest.mark.parametrize("antialiasing", [False, True])
def test_forward_state_vs_oracle(dgr, oracle, W, H, antialiasing):
    gs, s = small_scene(P=700, W=W, H=H, antialiasing=antialiasing, bg=(0.2, 0.5, 0.9), seed=W)
    want = run_oracle(oracle, gs, s)["fw"]
    got = ours_forward_state(gs, s)
    assert np.array_equal(got["radii"], want["radii"])
    assert got["num_rendered"] == want["R"]
    assert np.array_equal(got["point_list_keys"].view(np.uint64), want["keys"])
    assert np.array_equal(got["point_list"].view(np.uint32), want["point_list"])
    assert np.array_equal(got["ranges"].view(np.uint32), want["ranges"])
    assert np.abs(got["color"] - want["color"]).max() <= IMG_TOL
    assert np.abs(got["invdepth"] - want["invdepth"]).max() <= IMG_TOL
    assert np.mean(got["n_contrib"].view(np.uint32) != want["n_contrib"]) < 1e-3
    assert np.abs(got["final_T"] - want["final_T"]).max() <= IMG_TOL


@pytest.mark.parametrize("antialiasing", [False, True])
@pytest.mark.parametrize("bg", [(0.0, 0.0, 0.0), (1.0, 1.0, 1.0), (0.3, 0.6, 0.1)])
def test_backward_vs_oracle(dgr, oracle, antialiasing, bg):
    W, H = 80, 64
    gs, s = small_scene(P=900, W=W, H=H, antialiasing=antialiasing, bg=bg, seed=21)
    dc, dd = _upstream(W, H)
    want = run_oracle(oracle, gs, s, dc, dd)
    got = run_module(dgr, gs, s, dc, dd)
    assert np.array_equal(got["radii"], want["radii"])
    assert np.abs(got["color"] - want["color"]).max() <= IMG_TOL
    _check_grads(got["grads"], want["grads"], ("means3D", "means2D", "opacities", "shs", "scales", "rotations"))


def _ref_mean(ref, gs, s, dc, dd, runs=3):
    """Mean gradients of `runs` runs of the reference CUDA build (its atomics are order-nondeterministic)."""
    outs = [run_module(ref, gs, s, dc, dd) for _ in range(runs)]
    want = dict(outs[0])
    want["grads"] = {k: np.mean([o["grads"][k] for o in outs], axis=0) for k in outs[0]["grads"]}
    return want


def _checker(name, oracle, ref_restored):
    """'oracle' -> the C restatement on the CPU, 'reference' -> the reference's own CUDA build on this GPU."""
    if name == "oracle":
        return lambda gs, s, dc, dd: run_oracle(oracle, gs, s, dc, dd)
    return lambda gs, s, dc, dd: _ref_mean(ref_restored, gs, s, dc, dd)


CHECKERS = ["oracle", "reference"]


@pytest.mark.parametrize("checker", CHECKERS)
def test_backward_without_depth_gradient(dgr, oracle, ref_restored, checker):
    W, H = 64, 48
    gs, s = small_scene(P=500, W=W, H=H, bg=(0.5, 0.5, 0.5), seed=3)
    dc, _ = _upstream(W, H)
    want = _checker(checker, oracle, ref_restored)(gs, s, dc, None)
    got = run_module(dgr, gs, s, dc, None)
    _check_grads(got["grads"], want["grads"], ("means3D", "means2D", "opacities", "shs", "scales", "rotations"))


@pytest.mark.parametrize("checker", CHECKERS)
def test_precomputed_colour_and_covariance(dgr, oracle, ref_restored, checker):
    W, H, P = 72, 56, 600
    gs, s = small_scene(P=P, W=W, H=H, antialiasing=True, bg=(0.1, 0.2, 0.3), seed=9)
    from oracle import autograd_ref
    R = autograd_ref.quat_to_rot(gs["rotations"].double())
    Mx = R * gs["scales"].double()[:, None, :]
    Sig = Mx @ Mx.transpose(1, 2)
    cov6 = torch.stack([Sig[:, 0, 0], Sig[:, 0, 1], Sig[:, 0, 2], Sig[:, 1, 1], Sig[:, 1, 2], Sig[:, 2, 2]], 1)
    g = torch.Generator().manual_seed(4)
    gs2 = dict(means3D=gs["means3D"], opacities=gs["opacities"], colors_precomp=torch.rand(P, 3, generator=g),
               cov3D_precomp=cov6.float().contiguous())
    dc, dd = _upstream(W, H)
    want = _checker(checker, oracle, ref_restored)(gs2, s, dc, dd)
    got = run_module(dgr, gs2, s, dc, dd)
    assert np.array_equal(got["radii"], want["radii"])
    assert np.abs(got["color"] - want["color"]).max() <= IMG_TOL
    _check_grads(got["grads"], want["grads"], ("means3D", "means2D", "opacities", "colors_precomp", "cov3D_precomp"))


@pytest.mark.parametrize("checker", CHECKERS)
@pytest.mark.parametrize("deg,M", [(0, 1), (0, 16), (1, 4), (2, 9), (2, 16), (3, 16)])
def test_sh_degrees(dgr, oracle, ref_restored, deg, M, checker):
    W, H, P = 56, 40, 400
    gs, s = small_scene(P=P, W=W, H=H, sh_degree=3, seed=30 + deg)
    gs["shs"] = gs["shs"][:, :M, :].contiguous()
    s["sh_degree"] = deg
    dc, dd = _upstream(W, H)
    want = _checker(checker, oracle, ref_restored)(gs, s, dc, dd)
    got = run_module(dgr, gs, s, dc, dd)
    assert np.abs(got["color"] - want["color"]).max() <= IMG_TOL
    assert got["grads"]["shs"].shape == (P, M, 3)
    _check_grads(got["grads"], want["grads"], ("means3D", "opacities", "shs"))
    if M > (deg + 1) ** 2:
        assert np.all(got["grads"]["shs"][:, (deg + 1) ** 2:, :] == 0)


@pytest.mark.parametrize("P", [1, 31, 33, 257])
def test_ragged_counts(dgr, oracle, P):
    W, H = 33, 17
    gs, s = small_scene(P=P, W=W, H=H, seed=P, behind_fraction=0.0 if P == 1 else 0.2)
    dc, dd = _upstream(W, H)
    want = run_oracle(oracle, gs, s, dc, dd)
    got = run_module(dgr, gs, s, dc, dd)
    assert np.array_equal(got["radii"], want["radii"])
    assert np.abs(got["color"] - want["color"]).max() <= IMG_TOL
    _check_grads(got["grads"], want["grads"], ("means3D", "opacities", "shs", "scales", "rotations"))


def test_empty_input(dgr):
    gs, s = small_scene(P=8, W=32, H=24, bg=(0.2, 0.3, 0.4))
    gs = {k: v[:0].contiguous() for k, v in gs.items()}
    out = run_module(dgr, gs, s)
    # rasterize_points.cu:83-107: P == 0 => no launch, zero-filled outputs (no background term)
    assert out["color"].shape == (3, 24, 32) and np.all(out["color"] == 0)
    assert out["radii"].shape == (0,) and np.all(out["invdepth"] == 0)


def test_all_culled(dgr):
    W, H = 40, 40
    gs, s = small_scene(P=100, W=W, H=H, bg=(0.2, 0.3, 0.4))
    gs["means3D"][:, 2] = -50.0
    dc, dd = _upstream(W, H)
    out = run_module(dgr, gs, s, dc, dd)
    assert np.all(out["radii"] == 0) and np.all(out["invdepth"] == 0)
    assert np.allclose(out["color"][1], 0.3)
    assert all(np.all(v == 0) for v in out["grads"].values())


def test_mark_visible(dgr, oracle, ref_restored):
    from tests.helpers import settings_for
    gs, s = small_scene(P=5000, behind_fraction=0.4)
    pos = gs["means3D"].cuda()
    got = dgr.GaussianRasterizer(settings_for(dgr, s)).markVisible(pos)
    want = ref_restored.GaussianRasterizer(settings_for(ref_restored, s)).markVisible(pos)
    assert got.dtype == torch.bool and torch.equal(got, want)
    cpu = oracle.mark_visible(gs["means3D"].numpy(), s["viewmatrix"].numpy())
    assert (got.cpu().numpy() != cpu).sum() <= 2


def test_large_image_many_bands(dgr, oracle, ref_restored):
    """T = 257*257 = 66049 tiles: 9 bands of 31 tile rows in the tile-list construction (binning.cu); the reference
    sorts 17 tile bits here."""
    W = H = 4112
    gs, s = small_scene(P=3000, W=W, H=H, scale_mult=2.0, seed=12)
    want = run_oracle(oracle, gs, s)["fw"]
    got = ours_forward_state(gs, s)
    assert oracle.higher_msb(257 * 257) == 17
    assert got["num_rendered"] == want["R"]
    assert np.array_equal(got["point_list_keys"].view(np.uint64), want["keys"])
    assert np.array_equal(got["point_list"].view(np.uint32), want["point_list"])
    assert np.array_equal(got["ranges"].view(np.uint32), want["ranges"])
    # 17M pixels: the host libm expf and the device expf differ by an ulp on a few pairs, which
    # flips the alpha < 1/255 skip test there (a 4e-3 step).  Against the CPU oracle the image is
    # therefore checked up to a handful of such pixels; against the reference's CUDA build exactly.
    bad = np.abs(got["color"] - want["color"]).max(axis=0) > IMG_TOL
    assert bad.mean() < 1e-5 and np.abs(got["color"] - want["color"]).max() < 5e-3
    ref = ref_forward_state(ref_restored, gs, s)
    assert np.array_equal(got["point_list"], ref["point_list"])
    assert np.array_equal(got["ranges"], ref["ranges"])
    assert np.abs(got["color"] - ref["color"]).max() <= IMG_TOL
    assert np.array_equal(got["n_contrib"], ref["n_contrib"])


# ---------------------------------------------------------------- vs the reference's CUDA build
REF_CASES = [
    # P, W, H, scale_mult, antialiasing, bg
    (20000, 320, 240, 2.0, False, (0.0, 0.0, 0.0)),
    (20000, 333, 247, 3.0, True, (1.0, 1.0, 1.0)),
    (200000, 800, 800, 1.0, True, (0.2, 0.4, 0.7)),
    (100000, 1297, 840, 1.5, False, (0.3, 0.6, 0.1)),      # antialiasing off above 320x240, ragged right / bottom tiles
]


@pytest.mark.parametrize("P,W,H,scale_mult,aa,bg", REF_CASES)
def test_forward_state_bit_exact_vs_reference_cuda(dgr, ref_restored, P, W, H, scale_mult, aa, bg):
    gs, s = small_scene(P=P, W=W, H=H, scale_mult=scale_mult, antialiasing=aa, bg=bg, seed=P % 97)
    # exact depth ties exercise the stability of the ordering
    gs["means3D"][P // 2:P // 2 + P // 10] = gs["means3D"][:P // 10]
    want = ref_forward_state(ref_restored, gs, s)
    got = ours_forward_state(gs, s)
    assert got["num_rendered"] == want["num_rendered"]
    assert np.array_equal(got["radii"], want["radii"])
    assert np.array_equal(got["point_list_keys"], want["point_list_keys"])
    assert np.array_equal(got["point_list"], want["point_list"])
    assert np.array_equal(got["ranges"], want["ranges"])
    assert np.abs(got["color"] - want["color"]).max() <= IMG_TOL
    assert np.abs(got["invdepth"] - want["invdepth"]).max() <= IMG_TOL
    assert np.array_equal(got["n_contrib"], want["n_contrib"])
    assert np.abs(got["final_T"] - want["final_T"]).max() <= IMG_TOL


@pytest.mark.parametrize("P,W,H,scale_mult,aa,bg", REF_CASES)
def test_gradients_vs_reference_cuda(dgr, ref_restored, ref_asis, P, W, H, scale_mult, aa, bg):
    gs, s = small_scene(P=P, W=W, H=H, scale_mult=scale_mult, antialiasing=aa, bg=bg, seed=P % 97)
    dc, dd = _upstream(W, H)
    got = run_module(dgr, gs, s, dc, dd)
    # the reference accumulates with atomics in arbitrary order: compare against the mean of 3 runs
    runs = [run_module(ref_restored, gs, s, dc, dd) for _ in range(3)]
    want = {k: np.mean([r["grads"][k] for r in runs], axis=0) for k in runs[0]["grads"]}
    assert np.array_equal(got["radii"], runs[0]["radii"])
    assert np.abs(got["color"] - runs[0]["color"]).max() <= IMG_TOL
    _check_grads(got["grads"], want, ("means3D", "means2D", "opacities", "shs", "scales", "rotations"), outliers=1e-6)
    # the fork as shipped (truncated computeCov2DCUDA) is a valid oracle for these:
    asis = run_module(ref_asis, gs, s, dc, dd)["grads"]
    _check_grads(got["grads"], asis, ("means2D", "opacities", "shs"))


# ---------------------------------------------------------------- autograd behaviour
def test_retain_graph_and_accumulation(dgr):
    from tests.helpers import settings_for
    W, H = 64, 48
    gs, s = small_scene(P=800, W=W, H=H, seed=2)
    rs = settings_for(dgr, s)
    leaves = {k: v.cuda().requires_grad_(True) for k, v in gs.items()}
    m2 = torch.zeros(800, 3, device="cuda", requires_grad=True)
    color, radii, invd = dgr.GaussianRasterizer(rs)(
        means3D=leaves["means3D"], means2D=m2, opacities=leaves["opacities"], shs=leaves["shs"],
        scales=leaves["scales"], rotations=leaves["rotations"])
    assert color.dtype == torch.float32 and radii.dtype == torch.int32 and invd.shape == (1, H, W)
    assert not radii.requires_grad
    loss = color.sum() + invd.sum()
    loss.backward(retain_graph=True)
    g1 = leaves["shs"].grad.clone()
    loss.backward()
    assert rel_err(leaves["shs"].grad.cpu().numpy(), (2 * g1).cpu().numpy()) < 1e-5


def test_runs_on_side_stream_and_is_deterministic(dgr):
    W, H = 96, 64
    gs, s = small_scene(P=3000, W=W, H=H, seed=8)
    a = run_module(dgr, gs, s)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        b = run_module(dgr, gs, s)
    st.synchronize()
    assert np.array_equal(a["color"], b["color"]) and np.array_equal(a["radii"], b["radii"])
    assert np.array_equal(a["invdepth"], b["invdepth"])


def test_concurrent_streams_do_not_share_scratch(dgr):
    """Two rasterizations in flight on two streams (different scenes) must not disturb each other."""
    W, H = 160, 120
    scenes = [small_scene(P=20000, W=W, H=H, seed=31), small_scene(P=15000, W=W, H=H, seed=32, bg=(0.3, 0.3, 0.3))]
    want = [run_module(dgr, gs, s) for gs, s in scenes]
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    for rep in range(3):
        got = []
        for (gs, s), st in zip(scenes, streams):
            with torch.cuda.stream(st):
                got.append(run_module(dgr, gs, s))
        for st in streams:
            st.synchronize()
        for g, w in zip(got, want):
            assert np.array_equal(g["color"], w["color"]) and np.array_equal(g["radii"], w["radii"])


def test_debug_mode_runs_and_names_stages(dgr, oracle):
    """debug=True synchronises and checks after every stage (role of CHECK_CUDA, auxiliary.h:193-200)."""
    W, H = 64, 48
    gs, s = small_scene(P=500, W=W, H=H, seed=3)
    s["debug"] = True
    dc, dd = _upstream(W, H)
    want = run_oracle(oracle, gs, s, dc, dd)
    got = run_module(dgr, gs, s, dc, dd)
    assert np.abs(got["color"] - want["color"]).max() <= IMG_TOL
    _check_grads(got["grads"], want["grads"], ("means3D", "opacities", "shs", "scales", "rotations"))


def test_workspace_regrowth_keeps_results(dgr):
    """Growing then shrinking P re-uses / regrows the scratch arena (config C5 behaviour)."""
    from gaussian_splatting_b200 import rasterizer
    W, H = 128, 96
    outs = {}
    for P in (500, 20000, 700, 20000):
        gs, s = small_scene(P=P, W=W, H=H, seed=1)
        o = run_module(dgr, gs, s)
        if P in outs:
            assert np.array_equal(outs[P]["color"], o["color"])
        outs[P] = o
    st = rasterizer.workspace_stats()
    assert st["scratch_bytes"] > 0 and st["regrows"] >= 1


def test_accumulator_cleared_by_the_forward_survives_interleaving(dgr):
    """GSR_ACC_CLEAR_IN_FORWARD / GSR_ACC_IS_CLEAR (include/gsrast.h): the blend forward of a view that needs gradients
    zeroes the backward's accumulator; the backward may only rely on that when nothing used the arena in between.
    Forwards and backwards of views of different and of equal sizes interleaved on one stream must give the gradients
    of the views run one after the other."""
    from gaussian_splatting_b200 import rasterizer
    from tests.helpers import settings_for
    dev = torch.device("cuda", torch.cuda.current_device())
    views = [small_scene(P=20000, W=320, H=240, seed=41), small_scene(P=15000, W=200, H=152, seed=42),
             small_scene(P=20000, W=320, H=240, seed=43, bg=(0.2, 0.5, 0.1))]
    ups = [_upstream(s["image_width"], s["image_height"], seed=7 + i) for i, (_, s) in enumerate(views)]
    want = [run_module(dgr, gs, s, dc, dd)["grads"] for (gs, s), (dc, dd) in zip(views, ups)]
    keys = ("means3D", "means2D", "opacities", "shs", "scales", "rotations")

    def forward(i):
        gs, s = views[i]
        leaves = {k: v.cuda().clone().requires_grad_(True) for k, v in gs.items()}
        m2 = torch.zeros(gs["means3D"].shape[0], 3, device="cuda", requires_grad=True)
        color, radii, invd = dgr.GaussianRasterizer(settings_for(dgr, s))(
            means3D=leaves["means3D"], means2D=m2, opacities=leaves["opacities"], shs=leaves["shs"],
            scales=leaves["scales"], rotations=leaves["rotations"])
        dc, dd = ups[i]
        return leaves, m2, (color * dc.cuda()).sum() + (invd * dd.cuda()).sum()

    def grads(leaves, m2):
        g = {k: v.grad.cpu().numpy() for k, v in leaves.items()}
        g["means2D"] = m2.grad.cpu().numpy()
        return g

    for order in ([0, 1], [0, 2], [1, 0, 2], [2, 2]):
        fw = [forward(i) for i in order]
        P, s = views[order[-1]][0]["means3D"].shape[0], views[order[-1]][1]
        assert rasterizer._workspace(dev).acc_clear_for == (P, s["image_width"], s["image_height"])
        for i, (leaves, m2, loss) in reversed(list(zip(order, fw))):
            loss.backward()
            assert rasterizer._workspace(dev).acc_clear_for is None
            _check_grads(grads(leaves, m2), want[i], keys, tol=1e-4)
    # a forward without gradients in between (other size: its buffers overlay the zeros; same size: they do not)
    for other in (1, 2):
        leaves, m2, loss = forward(0)
        with torch.no_grad():
            run_module(dgr, *views[other])
        loss.backward()
        _check_grads(grads(leaves, m2), want[0], keys, tol=1e-4)
    # twice through the same graph: the second backward must clear the accumulator itself
    leaves, m2, loss = forward(0)
    loss.backward(retain_graph=True)
    loss.backward()
    g = grads(leaves, m2)
    _check_grads({k: 0.5 * v for k, v in g.items()}, want[0], keys, tol=1e-4)


# ---------------------------------------------------------------- multi-GPU (only where >= 2 GPUs are visible)
def test_fused_multicast_reduction_matches_nccl():
    """tests/multi_gpu_check.py under torchrun: in-kernel NVLink-multicast gradient reduction vs NCCL."""
    import os
    import subprocess
    import sys
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (run tests/multi_gpu_check.py under torchrun on a multi-GPU box)")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29541",
                        os.path.join(root, "tests", "multi_gpu_check.py")],
                       capture_output=True, text=True, timeout=600, cwd=root)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "fused multimem reduction == NCCL all-reduce" in r.stdout


# ---------------------------------------------------------------- BASELINE.json full size
def test_full_size_c3_vs_reference_cuda(dgr, ref_restored):
    """C3 of BASELINE.json (3M Gaussians, 1920x1080, antialiasing) against the reference's CUDA build:
    bit-exact binning state and n_contrib, image / depth <= 1e-5, gradients <= 1e-3 relative."""
    from gaussian_splatting_b200 import synthetic
    cfg = synthetic.CONFIGS["C3"]
    gs, cam = synthetic.make_scene(cfg)
    s = dict(cam)
    s.update(bg=torch.tensor([0.0, 0.0, 0.0]), scale_modifier=1.0, sh_degree=3, prefiltered=False, debug=False)
    want = ref_forward_state(ref_restored, gs, s)
    got = ours_forward_state(gs, s)
    assert got["num_rendered"] == want["num_rendered"] > 20_000_000
    assert np.array_equal(got["radii"], want["radii"])
    assert np.array_equal(got["point_list"], want["point_list"])
    assert np.array_equal(got["point_list_keys"], want["point_list_keys"])
    assert np.array_equal(got["ranges"], want["ranges"])
    assert np.array_equal(got["n_contrib"], want["n_contrib"])
    assert np.abs(got["color"] - want["color"]).max() <= IMG_TOL
    assert np.abs(got["invdepth"] - want["invdepth"]).max() <= IMG_TOL
    del want, got
    torch.cuda.empty_cache()
    dc, dd = _upstream(cfg.W, cfg.H)
    a = run_module(dgr, gs, s, dc, dd)
    b = _ref_mean(ref_restored, gs, s, dc, dd, runs=3)
    _check_grads(a["grads"], b["grads"], ("means3D", "means2D", "opacities", "shs", "scales", "rotations"), outliers=1e-6)


@pytest.mark.parametrize("name", ["C2", "C4"])
def test_full_size_c2_c4_vs_reference_cuda(dgr, ref_restored, name):
    """C2 (1M Gaussians, 1297x840: ragged right / bottom tile edge at scale) and C4 (6M Gaussians, 3840x2160:
    32 400 tiles, four bands of tile rows) of BASELINE.json against the reference's CUDA build: binning state and
    n_contrib bit-exact, image / depth <= 1e-5, gradients as _check_grads."""
    from gaussian_splatting_b200 import synthetic
    cfg = synthetic.CONFIGS[name]
    gs, cam = synthetic.make_scene(cfg)
    s = dict(cam)
    s.update(bg=torch.tensor([0.1, 0.2, 0.3]), scale_modifier=1.0, sh_degree=3, prefiltered=False, debug=False)
    want = ref_forward_state(ref_restored, gs, s)
    got = ours_forward_state(gs, s)
    assert got["num_rendered"] == want["num_rendered"] > 1_000_000
    assert np.array_equal(got["radii"], want["radii"])
    assert np.array_equal(got["point_list"], want["point_list"])
    assert np.array_equal(got["ranges"], want["ranges"])
    assert np.array_equal(got["n_contrib"], want["n_contrib"])
    assert np.abs(got["color"] - want["color"]).max() <= IMG_TOL
    assert np.abs(got["invdepth"] - want["invdepth"]).max() <= IMG_TOL
    del want, got
    torch.cuda.empty_cache()
    dc, dd = _upstream(cfg.W, cfg.H)
    a = run_module(dgr, gs, s, dc, dd)
    b = _ref_mean(ref_restored, gs, s, dc, dd, runs=3)
    _check_grads(a["grads"], b["grads"], ("means3D", "means2D", "opacities", "shs", "scales", "rotations"), outliers=1e-6)
