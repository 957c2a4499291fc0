"""Differential fuzzer (not collected by pytest): random models and thresholds through csrc/densify.cu on the CUDA
emulation (via gaussian_splatting_b200.densify) against oracle/densify_ref.py — structure and copied values exact.
    python tests/host_emul/fuzz_densify.py
Round-1 run: 366 configurations (P 1..3000, SH rows 0/3/8/15, gradient thresholds 0..1, extents 0.5..50, with / without
max_screen_size, min_opacity 0.005..0.5): 0 mismatches."""
import sys, time, types, contextlib; sys.path.insert(0, "/root/repo")
import numpy as np, torch
from torch import nn
import tests.test_densify_emul_cpu as E
from gaussian_splatting_b200 import _lib, densify
from oracle import densify_ref as dr
lib = E.load_kernel_emul()
_lib.load = lambda: lib
densify._require_cuda = lambda t, what: None
densify._device_ctx = lambda d: contextlib.nullcontext()
densify._stream = lambda d: None
densify._wait = lambda d: None
densify._pinned_counts = lambda: torch.zeros(4, dtype=torch.int32)
rng = np.random.default_rng(5)
t0 = time.time(); n = 0; fails = 0
while time.time() - t0 < 150:
    P = int(rng.integers(1, 3000)); K = int(rng.choice([0, 3, 8, 15]))
    s = {"xyz": rng.normal(size=(P, 3)), "f_dc": rng.normal(size=(P, 1, 3)), "f_rest": rng.normal(size=(P, K, 3)),
         "opacity": rng.normal(size=(P, 1)) * 2.5 - 2, "scaling": rng.normal(size=(P, 3)) * rng.choice([0.3, 0.9, 2.0]) - rng.choice([1.0, 3.0, 5.0]),
         "rotation": rng.normal(size=(P, 4))}
    s = {k: v.astype(np.float32) for k, v in s.items()}
    for g in dr.GROUPS:
        s["m_" + g] = rng.normal(size=s[g].shape).astype(np.float32); s["v_" + g] = rng.random(size=s[g].shape).astype(np.float32)
    s["denom"] = rng.integers(0, 4, size=(P, 1)).astype(np.float32)
    s["xyz_gradient_accum"] = (rng.random(size=(P, 1)) * 0.0006 * s["denom"]).astype(np.float32)
    s["max_radii2D"] = (rng.random(size=(P,)) * 50).astype(np.float32)
    noise = rng.normal(size=(2, P, 3)).astype(np.float32)
    thr = float(rng.choice([0.0, 0.0001, 0.0002, 0.0004, 1.0])); ext = float(rng.choice([0.5, 5.0, 50.0]))
    mso = rng.choice([None, 20, 0.5]); mino = float(rng.choice([0.005, 0.05, 0.5])); pd = float(rng.choice([0.01, 0.001]))
    want, info = dr.densify_and_prune(s, thr, mino, ext, mso, np.zeros(P, np.int32), pd, noise)
    gm = types.SimpleNamespace(percent_dense=pd); groups = []
    for g in dr.GROUPS:
        p = nn.Parameter(torch.from_numpy(s[g].copy())); setattr(gm, densify._ATTR[g], p); groups.append({"params": [p], "name": g})
    gm.optimizer = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
    for g in dr.GROUPS:
        gm.optimizer.state[getattr(gm, densify._ATTR[g])] = {"step": torch.tensor(1.0), "exp_avg": torch.from_numpy(s["m_" + g].copy()), "exp_avg_sq": torch.from_numpy(s["v_" + g].copy())}
    gm.xyz_gradient_accum = torch.from_numpy(s["xyz_gradient_accum"].copy()); gm.denom = torch.from_numpy(s["denom"].copy()); gm.max_radii2D = torch.from_numpy(s["max_radii2D"].copy())
    try:
        densify.densify_and_prune(gm, thr, mino, ext, mso, None, noise=torch.from_numpy(noise))
        for g in dr.GROUPS:
            p = getattr(gm, densify._ATTR[g]); st = gm.optimizer.state[p]
            assert tuple(p.shape) == want[g].shape, (g, tuple(p.shape), want[g].shape)
            if g in ("xyz", "scaling"): np.testing.assert_allclose(p.detach().numpy(), want[g], rtol=3e-6, atol=3e-6)
            else: assert np.array_equal(p.detach().numpy(), want[g]), g
            assert np.array_equal(st["exp_avg"].numpy(), want["m_" + g]) and np.array_equal(st["exp_avg_sq"].numpy(), want["v_" + g])
    except AssertionError as e:
        fails += 1; print("FAIL", dict(P=P, K=K, thr=thr, ext=ext, mso=mso, mino=mino, pd=pd), str(e)[:300].replace("\n", " "), flush=True)
    n += 1
print("configs", n, "fails", fails)
