"""GPU parity tests of SURVEY.md 8f rank 3 (density control, csrc/densify.cu through gaussian_splatting_b200.densify)
against the golden vectors of the reference's own GaussianModel (tests/golden/densify_golden.npz) and the numpy oracle
(oracle/densify_ref.py); at BASELINE size (3M Gaussians) through size-independent properties (provenance ids, segment
order, moment rules).  Structure is exact; copied values bit-identical; recomputed values (children's xyz / scaling,
reset opacity) within 2e-6."""
import os
import types

import numpy as np
import pytest
import torch
from torch import nn

from oracle import densify_ref as dr

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = np.load(os.path.join(HERE, "golden", "densify_golden.npz"))
DEV = "cuda:0"
FULL_P = 3_000_000          # BASELINE C3


def _densify():
    from gaussian_splatting_b200 import densify
    return densify


def model_from(state, percent_dense=0.01, with_state=True, step=3.0):
    d = _densify()
    gm = types.SimpleNamespace(percent_dense=percent_dense)
    groups = []
    for g in dr.GROUPS:
        p = nn.Parameter(torch.from_numpy(np.ascontiguousarray(state[g])).to(DEV))
        setattr(gm, d._ATTR[g], p)
        groups.append({"params": [p], "lr": 1e-3, "name": g})
    gm.optimizer = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
    if with_state:
        for g in dr.GROUPS:
            gm.optimizer.state[getattr(gm, d._ATTR[g])] = {
                "step": torch.tensor(step), "exp_avg": torch.from_numpy(np.ascontiguousarray(state["m_" + g])).to(DEV),
                "exp_avg_sq": torch.from_numpy(np.ascontiguousarray(state["v_" + g])).to(DEV)}
    for k in ("xyz_gradient_accum", "denom", "max_radii2D"):
        setattr(gm, k, torch.from_numpy(np.ascontiguousarray(state[k])).to(DEV))
    return gm


def gold_state(prefix):
    keys = [pre + g for g in dr.GROUPS for pre in ("", "m_", "v_")] + ["xyz_gradient_accum", "denom", "max_radii2D"]
    return {k: GOLD[prefix + k] for k in keys}


def assert_model(gm, want, recomputed=(), with_state=True):
    d = _densify()
    for g in dr.GROUPS:
        p = getattr(gm, d._ATTR[g])
        assert isinstance(p, nn.Parameter) and p.requires_grad and p.device.type == torch.device(DEV).type
        tensors = [("", p.detach())]
        if with_state:
            st = gm.optimizer.state[p]
            tensors += [("m_", st["exp_avg"]), ("v_", st["exp_avg_sq"])]
        for pre, t in tensors:
            w = want[pre + g]
            assert tuple(t.shape) == w.shape, (pre + g, tuple(t.shape), w.shape)
            if pre == "" and g in recomputed:
                np.testing.assert_allclose(t.cpu().numpy(), w, rtol=2e-6, atol=2e-6, err_msg=g)
            else:
                assert np.array_equal(t.cpu().numpy(), w), pre + g
    for k in ("xyz_gradient_accum", "denom", "max_radii2D"):
        assert np.array_equal(getattr(gm, k).cpu().numpy(), want[k]), k


@pytest.mark.parametrize("case", ("a", "b", "c", "d", "e"))
def test_densify_and_prune_matches_reference_golden(case):
    d = _densify()
    max_grad, min_opacity, extent, size_thr, pd = GOLD[f"{case}_args"]
    gm = model_from(gold_state(f"{case}_in_"), percent_dense=float(pd))
    info = d.densify_and_prune(gm, float(max_grad), float(min_opacity), float(extent),
                               float(size_thr) if size_thr else None, torch.from_numpy(GOLD[f"{case}_radii"]).to(DEV),
                               noise=torch.from_numpy(GOLD[f"{case}_noise"]).to(DEV))
    assert_model(gm, gold_state(f"{case}_out_"), recomputed=("xyz", "scaling"))
    assert info["P"] == info["kept"] + info["clones"] + 2 * info["split"] == gm._xyz.shape[0]
    assert len(gm.optimizer.state) == 6 and all(float(s["step"]) == 3.0 for s in gm.optimizer.state.values())


def test_reset_opacity_prune_points_and_statistics_match_reference_golden():
    d = _densify()
    gm = model_from(gold_state("a_out_"))
    d.reset_opacity(gm)
    assert_model(gm, gold_state("a_reset_"), recomputed=("opacity",))
    d.prune_points(gm, torch.from_numpy(GOLD["a_prune_mask"]).to(DEV))
    assert_model(gm, gold_state("a_pruned_"), recomputed=("opacity",))
    P = GOLD["a_in_xyz"].shape[0]
    st = types.SimpleNamespace(xyz_gradient_accum=torch.zeros(P, 1, device=DEV), denom=torch.zeros(P, 1, device=DEV),
                               max_radii2D=torch.zeros(P, device=DEV))
    for v in range(4):
        vsp = types.SimpleNamespace(grad=torch.from_numpy(GOLD[f"a_view{v}_grad"]).to(DEV))
        vis = torch.from_numpy(GOLD[f"a_view{v}_vis"]).to(DEV)
        # boolean mask and render()'s `(radii > 0).nonzero()` index tensor (gaussian_renderer/__init__.py:152) alternate
        d.add_densification_stats(st, vsp, vis if v % 2 else vis.nonzero(),
                                  radii=torch.from_numpy(GOLD[f"a_view{v}_radii"]).to(DEV))
    np.testing.assert_allclose(st.xyz_gradient_accum.cpu().numpy(), GOLD["a_in_xyz_gradient_accum"], rtol=3e-7, atol=0)
    assert np.array_equal(st.denom.cpu().numpy(), GOLD["a_in_denom"])
    assert np.array_equal(st.max_radii2D.cpu().numpy(), GOLD["a_in_max_radii2D"])


def random_state(P, K, seed, ids=False):
    rng = np.random.default_rng(seed)
    s = {"xyz": rng.normal(size=(P, 3)), "f_dc": rng.normal(size=(P, 1, 3)), "f_rest": rng.normal(size=(P, K, 3)),
         "opacity": rng.normal(size=(P, 1)) * 2.5 - 2, "scaling": rng.normal(size=(P, 3)) * 0.9 - 3,
         "rotation": rng.normal(size=(P, 4))}
    s = {k: v.astype(np.float32) for k, v in s.items()}
    if ids:
        s["f_dc"][:, 0, 0] = np.arange(P, dtype=np.float32)          # provenance channel (exact below 2^24)
    for g in dr.GROUPS:
        s["m_" + g] = rng.normal(size=s[g].shape).astype(np.float32)
        s["v_" + g] = rng.random(size=s[g].shape).astype(np.float32)
    s["denom"] = rng.integers(0, 4, size=(P, 1)).astype(np.float32)
    s["xyz_gradient_accum"] = (rng.random(size=(P, 1)) * 0.0006 * s["denom"]).astype(np.float32)
    s["max_radii2D"] = (rng.random(size=(P,)) * 50).astype(np.float32)
    return s, rng.normal(size=(2, P, 3)).astype(np.float32)


@pytest.mark.parametrize("P,K", ((1, 15), (255, 0), (1024, 3), (1025, 8), (8000, 15)))
def test_densify_and_prune_matches_oracle(P, K):
    d = _densify()
    s, noise = random_state(P, K, seed=P)
    want, _ = dr.densify_and_prune(s, 0.0002, 0.005, 5.0, 20, np.zeros(P, np.int32), 0.01, noise)
    gm = model_from(s)
    d.densify_and_prune(gm, 0.0002, 0.005, 5.0, 20, None, noise=torch.from_numpy(noise).to(DEV))
    assert_model(gm, want, recomputed=("xyz", "scaling"))


def test_densify_before_the_first_optimizer_step_and_everything_pruned():
    d = _densify()
    s, noise = random_state(3000, 3, seed=11)
    want, _ = dr.densify_and_prune(s, 0.0002, 0.005, 5.0, None, np.zeros(3000, np.int32), 0.01, noise)
    gm = model_from(s, with_state=False)
    d.densify_and_prune(gm, 0.0002, 0.005, 5.0, None, None, noise=torch.from_numpy(noise).to(DEV))
    assert_model(gm, want, recomputed=("xyz", "scaling"), with_state=False)
    assert len(gm.optimizer.state) == 0
    # min_opacity above every opacity: the model becomes empty, like the reference's all-true prune mask
    gm = model_from(s)
    info = d.densify_and_prune(gm, 0.0002, 2.0, 5.0, 20, None, noise=torch.from_numpy(noise).to(DEV))
    assert info["P"] == 0 and gm._xyz.shape == (0, 3) and gm._features_rest.shape == (0, 3, 3)
    assert gm.optimizer.state[gm._xyz]["exp_avg"].shape == (0, 3) and gm.denom.shape == (0, 1)
    assert d.densify_and_prune(gm, 0.0002, 0.005, 5.0, 20, None)["P"] == 0           # empty model: no launch


def test_default_noise_is_seeded_and_statistically_right():
    d = _densify()
    s, _ = random_state(20000, 0, seed=5)
    out = []
    for _ in range(2):
        gm = model_from(s)
        gen = torch.Generator(device=DEV).manual_seed(123)
        info = d.densify_and_prune(gm, 0.0002, 0.005, 5.0, None, None, generator=gen)
        out.append(gm._xyz.detach().clone())
    assert torch.equal(out[0], out[1])               # same generator seed => identical replicas (view-parallel ranks)
    C, A, B = info["split"], info["kept"], info["clones"]
    assert C > 100
    # children are drawn around their parent: offsets of the two copies are independent with zero mean
    c0, c1 = out[0][A + B:A + B + C], out[0][A + B + C:]
    assert not torch.equal(c0, c1) and float((c0 - c1).mean().abs()) < 0.05


def test_full_size_properties():
    """P = 3M (BASELINE C3), SH degree 3, both moments: provenance, segment order and moment rules."""
    d = _densify()
    P, K = FULL_P, 15
    g = torch.Generator(device=DEV).manual_seed(0)
    r = lambda *shape: torch.randn(*shape, device=DEV, generator=g)
    s = {"xyz": r(P, 3), "f_dc": r(P, 1, 3), "f_rest": r(P, K, 3), "opacity": r(P, 1) * 2.5 - 2,
         "scaling": r(P, 3) * 0.9 - 3, "rotation": r(P, 4)}
    s["f_dc"][:, 0, 0] = torch.arange(P, device=DEV, dtype=torch.float32)
    gm = types.SimpleNamespace(percent_dense=0.01)
    groups = []
    for name in dr.GROUPS:
        p = nn.Parameter(s[name])
        setattr(gm, d._ATTR[name], p)
        groups.append({"params": [p], "name": name})
    gm.optimizer = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
    old = {}
    for name in dr.GROUPS:
        p = getattr(gm, d._ATTR[name])
        m, v = torch.randn_like(p), torch.rand_like(p)
        gm.optimizer.state[p] = {"step": torch.tensor(7.0), "exp_avg": m, "exp_avg_sq": v}
        old[name] = (p.detach(), m, v)
    gm.denom = torch.randint(0, 4, (P, 1), device=DEV, generator=g).float()
    gm.xyz_gradient_accum = torch.rand(P, 1, device=DEV, generator=g) * 0.0006 * gm.denom
    gm.max_radii2D = torch.rand(P, device=DEV, generator=g) * 50
    noise = r(2, P, 3)
    grads = torch.nan_to_num(gm.xyz_gradient_accum / gm.denom, nan=0.0)[:, 0]
    scal = torch.exp(old["scaling"][0])
    big = scal.max(1).values > 0.01 * 5.0
    info = d.densify_and_prune(gm, 0.0002, 0.005, 5.0, 20, None, noise=noise)
    A, B, C, Pn = info["kept"], info["clones"], info["split"], info["P"]
    assert Pn == A + B + 2 * C == gm._xyz.shape[0] and min(A, B, C) > 1000
    ids = gm._features_dc.detach()[:, 0, 0].long()
    segA, segB, segC0, segC1 = ids[:A], ids[A:A + B], ids[A + B:A + B + C], ids[A + B + C:]
    for seg in (segA, segB, segC0):
        assert bool((seg[1:] > seg[:-1]).all())                  # source order inside every segment
    assert torch.equal(segC0, segC1)
    # decisions agree with the reference's masks evaluated by torch on the same device
    # (a 1-ulp difference between torch's and the kernel's exp / sigmoid may flip a row that sits on a threshold)
    few = lambda bad: int(bad.sum()) <= 2
    sel = grads >= 0.0002
    assert few(~(sel & big)[segC0]) and few(~(sel & ~big)[segB]) and few((sel & big)[segA])
    low = torch.sigmoid(old["opacity"][0][:, 0]) < 0.005
    assert few(low[ids])
    assert few((scal.max(1).values > 0.5)[torch.cat((segA, segB))])
    # copied tensors and the moment rules (:462-466)
    for name in dr.GROUPS:
        p = getattr(gm, d._ATTR[name])
        st = gm.optimizer.state[p]
        src, m, v = old[name]
        assert torch.equal(st["exp_avg"][:A], m[segA]) and torch.equal(st["exp_avg_sq"][:A], v[segA])
        assert not bool(st["exp_avg"][A:].any()) and not bool(st["exp_avg_sq"][A:].any())
        assert float(st["step"]) == 7.0
        if name == "scaling":
            assert torch.equal(p.detach()[:A + B], src[ids[:A + B]])
            want = torch.log(scal[segC0] * 0.625)
            assert torch.allclose(p.detach()[A + B:A + B + C], want, rtol=2e-6, atol=2e-6)
            assert torch.equal(p.detach()[A + B:A + B + C], p.detach()[A + B + C:])
        elif name == "xyz":
            assert torch.equal(p.detach()[:A + B], src[ids[:A + B]])
            q = torch.nn.functional.normalize(old["rotation"][0][segC0])
            w, x, y, z = q.unbind(1)
            R = torch.stack((1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                             2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                             2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)), 1).reshape(-1, 3, 3)
            for copy, lo in ((0, A + B), (1, A + B + C)):
                want = torch.bmm(R, (scal[segC0] * noise[copy][segC0]).unsqueeze(-1)).squeeze(-1) + src[segC0]
                assert torch.allclose(p.detach()[lo:lo + C], want, rtol=1e-5, atol=1e-5)
        else:
            assert torch.equal(p.detach(), src[ids])
    assert not bool(gm.xyz_gradient_accum.any()) and gm.denom.shape == (Pn, 1) and gm.max_radii2D.shape == (Pn,)
    # prune_points is idempotent on an all-false mask and exact on a random one
    before = gm._features_rest.detach().clone()
    d.prune_points(gm, torch.zeros(Pn, dtype=torch.bool, device=DEV))
    assert torch.equal(gm._features_rest.detach(), before)
    mask = torch.rand(Pn, device=DEV, generator=g) < 0.4
    d.prune_points(gm, mask)
    assert torch.equal(gm._features_rest.detach(), before[~mask]) and gm.denom.shape == (int((~mask).sum()), 1)


def test_errors_are_loud():
    d = _densify()
    s, noise = random_state(100, 0, seed=1)
    gm = model_from(s)
    with pytest.raises(RuntimeError):
        d.prune_points(gm, torch.zeros(99, dtype=torch.bool, device=DEV))
    with pytest.raises(RuntimeError):
        d.densify_and_prune(gm, 0.0, 0.005, 5.0, 20, None, noise=torch.zeros(2, 50, 3, device=DEV))
    cpu = types.SimpleNamespace(percent_dense=0.01)
    p = nn.Parameter(torch.zeros(4, 3))
    cpu.optimizer = torch.optim.Adam([{"params": [p], "name": n} for n in ("xyz",)], lr=0.0)
    with pytest.raises(RuntimeError):
        d.densify_and_prune(cpu, 0.0002, 0.005, 5.0, 20, None)


def test_training_slice_render_stats_step_densify_save_load(tmp_path):
    """The callers either side of the rasterizer in one piece (train.py:143-251 without datasets): render -> L1 ->
    backward -> statistics from render()'s visibility filter -> SparseGaussianAdam -> densify_and_prune -> render with
    the new count -> save_ply -> load_ply -> identical image."""
    import diff_gaussian_rasterization as dgr
    from gaussian_splatting_b200 import plyio
    from tests.helpers import settings_for, small_scene
    d = _densify()
    W, H = 128, 96
    gs, s = small_scene(P=4000, W=W, H=H, seed=3, sh_degree=3, scale_mult=2.0, behind_fraction=0.05)
    gm = types.SimpleNamespace(percent_dense=0.01, max_sh_degree=3, active_sh_degree=3)
    raw = {"xyz": gs["means3D"], "f_dc": gs["shs"][:, :1], "f_rest": gs["shs"][:, 1:],
           "opacity": torch.logit(gs["opacities"].clamp(1e-4, 1 - 1e-4)), "scaling": torch.log(gs["scales"]),
           "rotation": gs["rotations"]}
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 1.25e-4, "opacity": 0.025, "scaling": 0.005, "rotation": 0.001}
    groups = []
    for name in dr.GROUPS:
        p = nn.Parameter(raw[name].contiguous().to(DEV))
        setattr(gm, d._ATTR[name], p)
        groups.append({"params": [p], "lr": lrs[name], "name": name})
    gm.optimizer = dgr.SparseGaussianAdam(groups, lr=0.0, eps=1e-15)
    P = gm._xyz.shape[0]
    gm.xyz_gradient_accum = torch.zeros(P, 1, device=DEV)
    gm.denom = torch.zeros(P, 1, device=DEV)
    gm.max_radii2D = torch.zeros(P, device=DEV)
    target = torch.rand(3, H, W, generator=torch.Generator().manual_seed(0)).to(DEV)
    rast = dgr.GaussianRasterizer(settings_for(dgr, s, DEV))

    def render(model):
        means2D = torch.zeros_like(model._xyz, requires_grad=True)          # screenspace_points (render():36-40)
        color, radii, invd = rast(means3D=model._xyz, means2D=means2D, opacities=torch.sigmoid(model._opacity),
                                  dc=model._features_dc, shs=model._features_rest, scales=torch.exp(model._scaling),
                                  rotations=torch.nn.functional.normalize(model._rotation))
        return color, radii, means2D

    for _ in range(3):
        color, radii, means2D = render(gm)
        (color - target).abs().mean().backward()
        d.add_densification_stats(gm, means2D, (radii > 0).nonzero(), radii=radii)
        gm.optimizer.step(radii > 0, radii.shape[0])
        gm.optimizer.zero_grad(set_to_none=True)
    vis = radii > 0
    assert int(vis.sum()) > 1000 and float(gm.denom.max()) == 3.0
    assert torch.equal(gm.max_radii2D[vis] >= radii[vis].float(), torch.ones_like(vis[vis]))
    # thresholds from the data so that every branch is taken
    g = torch.nan_to_num(gm.xyz_gradient_accum / gm.denom, nan=0.0)[:, 0]
    max_grad = float(g[g > 0].median())
    extent = float(torch.exp(gm._scaling.detach()).max(1).values.median()) / 0.01
    info = d.densify_and_prune(gm, max_grad, 0.005, extent, 20, radii)
    Pn = info["P"]
    assert info["clones"] > 0 and info["split"] > 0 and Pn != P and gm._xyz.shape[0] == Pn
    assert gm.xyz_gradient_accum.shape == (Pn, 1) and gm.max_radii2D.shape == (Pn,)
    # the new model renders, trains on (optimizer state was re-keyed and resized) and reports statistics
    color, radii, means2D = render(gm)
    assert color.shape == (3, H, W) and radii.shape == (Pn,) and bool(torch.isfinite(color).all())
    (color - target).abs().mean().backward()
    before = gm._xyz.detach().clone()
    gm.optimizer.step(radii > 0, Pn)
    assert not torch.equal(gm._xyz.detach(), before)
    assert gm.optimizer.state[gm._xyz]["exp_avg"].shape == (Pn, 3) and len(gm.optimizer.state) == 6
    d.add_densification_stats(gm, means2D, (radii > 0).nonzero(), radii=radii)
    assert float(gm.denom.sum()) == float((radii > 0).sum())
    gm.optimizer.zero_grad(set_to_none=True)
    d.reset_opacity(gm)
    assert float(torch.sigmoid(gm._opacity).max()) <= 0.0100001
    # file round trip: the loaded model renders the same image bit for bit
    with torch.no_grad():
        want, _, _ = render(gm)
    path = str(tmp_path / "point_cloud.ply")
    plyio.save_ply(gm, path)
    loaded = types.SimpleNamespace(max_sh_degree=3)
    plyio.load_ply(loaded, path, device=DEV)
    with torch.no_grad():
        got, _, _ = render(loaded)
    assert torch.equal(got, want)
