"""Kernel-logic check of the density-control path in the GPU-less container, with two emulation back-ends:

  "per-element"  csrc/densify_core.cuh (the per-element functions of csrc/densify.cu) compiled with g++ by
                 tests/host_emul/densify_emul.cpp, block scans / grid-stride loops replaced by serial loops;
  "kernels"      csrc/densify.cu ITSELF compiled as C++ against tests/host_emul/cuda_emul (a stand-in cuda_runtime.h and
                 a fiber scheduler: every CUDA thread of a block is a fiber, __syncthreads / __shfl_*_sync are scheduling
                 points): the real kernels, block scans, launch configurations and C-ABI argument checks run on the CPU.

The product's host code (gaussian_splatting_b200/densify.py, plyio.py) is driven on CPU tensors with its native calls
redirected to the emulation (monkeypatched in this test only — the product has no CPU path), and the result is compared
with the golden vectors of the reference's own GaussianModel (tests/golden/densify_golden.npz) and with the oracle.
The -m gpu tests (tests/test_zz_densify_gpu.py) run the same kernels on the device against the same vectors."""
import contextlib
import ctypes as C
import os
import subprocess
import types

import numpy as np
import pytest
import torch
from torch import nn

from gaussian_splatting_b200 import _lib, densify
from oracle import densify_ref as dr

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = np.load(os.path.join(HERE, "golden", "densify_golden.npz"))
ATTR = densify._ATTR
VP = C.c_void_p


def load_emul():
    """Builds (if stale) and binds tests/host_emul/libdensify_emul.so."""
    src = os.path.join(HERE, "host_emul", "densify_emul.cpp")
    out = os.path.join(HERE, "host_emul", "libdensify_emul.so")
    if not os.path.exists(out) or os.path.getmtime(out) < max(
            os.path.getmtime(src), os.path.getmtime(os.path.join(HERE, "..", "gaussian_splatting_b200", "csrc",
                                                                 "densify_core.cuh"))):
        subprocess.check_call(["g++", "-O2", "-shared", "-fPIC", "-ffp-contract=off", "-o", out, src])
    lib = C.CDLL(out)
    lib.emul_densify_plan.argtypes = [C.POINTER(_lib.GsrDensifyPlanArgs), VP, VP, VP, VP]
    lib.emul_densify_apply.argtypes = [C.POINTER(_lib.GsrDensifyApplyArgs), VP, VP, VP]
    lib.emul_densify_stats.argtypes = [C.c_int32, VP, VP, VP, C.c_int64] + [VP] * 4
    lib.emul_reset_opacity.argtypes = [C.c_int32] + [VP] * 3
    lib.emul_ply_rows.argtypes = [C.c_int, C.c_int32, C.c_int32] + [VP] * 7
    return lib


def load_kernel_emul(defines=(), tag=""):
    """Builds (if stale) tests/host_emul/libgsrast_emul.so (or a variant with extra defines): the product's .cu sources compiled as C++ (-DGSR_HOST_EMUL)
    against the CUDA stand-in of tests/host_emul/cuda_emul; exports the same C-ABI as libgsrast.so for those files."""
    emu_dir = os.path.join(HERE, "host_emul", "cuda_emul")
    csrc = os.path.join(HERE, "..", "gaussian_splatting_b200", "csrc")
    out = os.path.join(HERE, "host_emul", f"libgsrast{tag}_emul.so")
    units = [os.path.join(csrc, "densify.cu")]
    deps = units + [os.path.join(emu_dir, f) for f in os.listdir(emu_dir)] + \
        [os.path.join(csrc, f) for f in ("common.cuh", "block_scan.cuh", "densify_core.cuh")]
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(d) for d in deps):
        cmd = ["g++", "-std=c++17", "-O1", "-shared", "-fPIC", "-ffp-contract=off", "-U_FORTIFY_SOURCE", "-D_FORTIFY_SOURCE=0", "-DGSR_HOST_EMUL", "-I" + emu_dir]
        cmd += [f"-D{d}" for d in defines]
        for u in units:
            cmd += ["-x", "c++", u]
        cmd += ["-x", "c++", os.path.join(emu_dir, "emul.cpp"), os.path.join(emu_dir, "api_stub.cpp"), "-o", out]
        subprocess.check_call(cmd)
    lib = _lib.bind_densify(C.CDLL(out))
    lib.gsr_last_error.restype = C.c_char_p
    return lib


@pytest.fixture(scope="module")
def emul():
    return load_emul()


@pytest.fixture(scope="module")
def kernel_emul():
    return load_kernel_emul()


class FakeLib:
    """The C-ABI entry points of csrc/densify.cu, served by the CPU emulation (same structs, host pointers)."""

    def __init__(self, emul):
        self.e = emul

    def gsr_densify_plan_words(self, P):
        return 3 * P + 8

    def gsr_densify_plan(self, args, plan, counts, stream):
        P = args._obj.P
        return self.e.emul_densify_plan(args, plan, plan + 4 * P, plan + 8 * P, counts)

    def gsr_densify_apply(self, args, plan, stream):
        P = args._obj.P
        return self.e.emul_densify_apply(args, plan, plan + 4 * P, plan + 8 * P)

    def gsr_densify_stats(self, P, grad, vis, idx, n_idx, radii, accum, denom, maxr, stream):
        return self.e.emul_densify_stats(P, grad, vis, idx, n_idx, radii, accum, denom, maxr)

    def gsr_reset_opacity(self, P, op, m, v, stream):
        return self.e.emul_reset_opacity(P, op, m, v)

    def gsr_ply_pack(self, P, K, xyz, f_dc, f_rest, op, sc, rot, rows, stream):
        return self.e.emul_ply_rows(1, P, K, xyz, f_dc, f_rest, op, sc, rot, rows)

    def gsr_ply_unpack(self, P, K, rows, xyz, f_dc, f_rest, op, sc, rot, stream):
        return self.e.emul_ply_rows(0, P, K, xyz, f_dc, f_rest, op, sc, rot, rows)

    def gsr_last_error(self):
        return b"emulation"


@pytest.fixture(scope="module")
def kernel_emul_wide():
    # the 64-bit index instantiations of the move / PLY kernels (taken above 2^32 elements) forced at test sizes
    return load_kernel_emul(defines=("GSR_INDEX32_LIMIT=0",), tag="_wide")


@pytest.fixture(params=("per-element", "kernels", "kernels-64bit-index"))
def on_cpu(request, emul, kernel_emul, kernel_emul_wide, monkeypatch):
    fake = {"per-element": FakeLib(emul), "kernels": kernel_emul, "kernels-64bit-index": kernel_emul_wide}[request.param]
    monkeypatch.setattr(_lib, "load", lambda: fake)
    for mod in (densify,):
        monkeypatch.setattr(mod, "_require_cuda", lambda t, what: None)
        monkeypatch.setattr(mod, "_device_ctx", lambda d: contextlib.nullcontext())
        monkeypatch.setattr(mod, "_stream", lambda d: None)
        monkeypatch.setattr(mod, "_wait", lambda d: None)
        monkeypatch.setattr(mod, "_pinned_counts", lambda: torch.zeros(4, dtype=torch.int32))
    return fake


def make_model(prefix, with_state=True):
    """A stand-in with the attribute / optimizer-group layout of the reference's GaussianModel (:228-245)."""
    gm = types.SimpleNamespace()
    groups = []
    for g in dr.GROUPS:
        p = nn.Parameter(torch.from_numpy(GOLD[f"{prefix}{g}"].copy()))
        setattr(gm, ATTR[g], p)
        groups.append({"params": [p], "lr": 1e-3, "name": g})
    gm.optimizer = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
    if with_state:
        for g in dr.GROUPS:
            p = getattr(gm, ATTR[g])
            gm.optimizer.state[p] = {"step": torch.tensor(3.0),
                                     "exp_avg": torch.from_numpy(GOLD[f"{prefix}m_{g}"].copy()),
                                     "exp_avg_sq": torch.from_numpy(GOLD[f"{prefix}v_{g}"].copy())}
    gm.xyz_gradient_accum = torch.from_numpy(GOLD[f"{prefix}xyz_gradient_accum"].copy())
    gm.denom = torch.from_numpy(GOLD[f"{prefix}denom"].copy())
    gm.max_radii2D = torch.from_numpy(GOLD[f"{prefix}max_radii2D"].copy())
    gm.percent_dense = 0.01
    return gm


def assert_model(gm, prefix, recomputed=(), with_state=True):
    names = {g["name"]: g for g in gm.optimizer.param_groups}
    for g in dr.GROUPS:
        p = getattr(gm, ATTR[g])
        assert isinstance(p, nn.Parameter) and p.requires_grad and names[g]["params"][0] is p
        tensors = [("", p.detach())]
        if with_state:
            st = gm.optimizer.state[p]
            assert float(st["step"]) == 3.0 and len(gm.optimizer.state) == 6
            tensors += [("m_", st["exp_avg"]), ("v_", st["exp_avg_sq"])]
        for pre, t in tensors:
            want = GOLD[f"{prefix}{pre}{g}"]
            assert tuple(t.shape) == want.shape, (pre + g, t.shape, want.shape)
            if pre == "" and g in recomputed:
                np.testing.assert_allclose(t.numpy(), want, rtol=2e-6, atol=2e-6, err_msg=g)
            else:
                assert np.array_equal(t.numpy(), want), pre + g
    for k in ("xyz_gradient_accum", "denom", "max_radii2D"):
        assert np.array_equal(getattr(gm, k).numpy(), GOLD[prefix + k]), k


@pytest.mark.parametrize("case", ("a", "b", "c", "d", "e"))
def test_densify_and_prune_matches_reference(on_cpu, case):
    gm = make_model(f"{case}_in_")
    max_grad, min_opacity, extent, size_thr, pd = GOLD[f"{case}_args"]
    gm.percent_dense = float(pd)
    info = densify.densify_and_prune(gm, float(max_grad), float(min_opacity), float(extent),
                                     float(size_thr) if size_thr else None, torch.from_numpy(GOLD[f"{case}_radii"]),
                                     noise=torch.from_numpy(GOLD[f"{case}_noise"].copy()))
    assert_model(gm, f"{case}_out_", recomputed=("xyz", "scaling"))
    assert info["P"] == GOLD[f"{case}_out_xyz"].shape[0] == info["kept"] + info["clones"] + 2 * info["split"]
    assert gm.tmp_radii is None


def test_densify_without_optimizer_state(on_cpu):
    # before the first optimizer step the reference's state.get() returns None (:417,:426-428,:458,:478-481)
    gm = make_model("a_in_", with_state=False)
    max_grad, min_opacity, extent, size_thr, pd = GOLD["a_args"]
    densify.densify_and_prune(gm, float(max_grad), float(min_opacity), float(extent), float(size_thr), None,
                              noise=torch.from_numpy(GOLD["a_noise"].copy()))
    assert_model(gm, "a_out_", recomputed=("xyz", "scaling"), with_state=False)
    assert len(gm.optimizer.state) == 0


def test_reset_opacity_and_prune_points_match_reference(on_cpu):
    gm = make_model("a_out_")
    old = gm._opacity
    densify.reset_opacity(gm)
    assert gm._opacity is not old and old not in gm.optimizer.state
    assert_model(gm, "a_reset_", recomputed=("opacity",))
    gm.tmp_radii = torch.arange(gm._xyz.shape[0], dtype=torch.float32)
    mask = torch.from_numpy(GOLD["a_prune_mask"])
    densify.prune_points(gm, mask)
    assert_model(gm, "a_pruned_", recomputed=("opacity",))
    assert torch.equal(gm.tmp_radii, torch.arange(mask.numel(), dtype=torch.float32)[~mask])


@pytest.mark.parametrize("case", ("a", "c"))
@pytest.mark.parametrize("filter_kind", ("mask", "nonzero", "radii"))
def test_statistics_match_reference(on_cpu, case, filter_kind):
    P = GOLD[f"{case}_in_xyz"].shape[0]
    gm = types.SimpleNamespace(xyz_gradient_accum=torch.zeros(P, 1), denom=torch.zeros(P, 1), max_radii2D=torch.zeros(P))
    for v in range(4):
        vsp = types.SimpleNamespace(grad=torch.from_numpy(GOLD[f"{case}_view{v}_grad"].copy()))
        vis = torch.from_numpy(GOLD[f"{case}_view{v}_vis"])
        radii = torch.from_numpy(GOLD[f"{case}_view{v}_radii"])
        if filter_kind == "radii":
            # the golden filter is a random 70 % mask; make radii carry it (render(): visible <=> radii > 0)
            radii = torch.where(vis, radii.clamp_min(1), torch.zeros_like(radii))
            flt = None
        else:
            flt = vis if filter_kind == "mask" else vis.nonzero()       # render() returns the (V, 1) index tensor
        densify.add_densification_stats(gm, vsp, flt, radii=radii)
    if filter_kind == "radii":
        assert np.array_equal(gm.denom.numpy(), GOLD[f"{case}_in_denom"])
        return
    np.testing.assert_allclose(gm.xyz_gradient_accum.numpy(), GOLD[f"{case}_in_xyz_gradient_accum"], rtol=3e-7, atol=0)
    assert np.array_equal(gm.denom.numpy(), GOLD[f"{case}_in_denom"])
    assert np.array_equal(gm.max_radii2D.numpy(), GOLD[f"{case}_in_max_radii2D"])


def test_random_models_against_the_oracle(on_cpu):
    """Sizes that cross the plan kernels' block boundary (1024 Gaussians), against oracle/densify_ref.py."""
    rng = np.random.default_rng(7)
    for P, K in ((1023, 0), (1025, 3), (5000, 15)):
        s = {"xyz": rng.normal(size=(P, 3)), "f_dc": rng.normal(size=(P, 1, 3)), "f_rest": rng.normal(size=(P, K, 3)),
             "opacity": rng.normal(size=(P, 1)) * 2.5 - 2, "scaling": rng.normal(size=(P, 3)) * 0.9 - 3,
             "rotation": rng.normal(size=(P, 4))}
        s = {k: v.astype(np.float32) for k, v in s.items()}
        for g in dr.GROUPS:
            s["m_" + g] = rng.normal(size=s[g].shape).astype(np.float32)
            s["v_" + g] = rng.random(size=s[g].shape).astype(np.float32)
        s["denom"] = rng.integers(0, 4, size=(P, 1)).astype(np.float32)
        s["xyz_gradient_accum"] = (rng.random(size=(P, 1)) * 0.0006 * s["denom"]).astype(np.float32)
        s["max_radii2D"] = rng.random(size=(P,)).astype(np.float32) * 50
        noise = rng.normal(size=(2, P, 3)).astype(np.float32)
        want, info = dr.densify_and_prune(s, 0.0002, 0.005, 5.0, 20, np.zeros(P, np.int32), 0.01, noise)
        assert info["n_clone"] > 0 and info["n_split"] > 0
        gm = types.SimpleNamespace(percent_dense=0.01)
        groups = []
        for g in dr.GROUPS:
            p = nn.Parameter(torch.from_numpy(s[g].copy()))
            setattr(gm, ATTR[g], p)
            groups.append({"params": [p], "name": g})
        gm.optimizer = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
        for g in dr.GROUPS:
            gm.optimizer.state[getattr(gm, ATTR[g])] = {"step": torch.tensor(1.0),
                                                        "exp_avg": torch.from_numpy(s["m_" + g].copy()),
                                                        "exp_avg_sq": torch.from_numpy(s["v_" + g].copy())}
        gm.xyz_gradient_accum = torch.from_numpy(s["xyz_gradient_accum"].copy())
        gm.denom = torch.from_numpy(s["denom"].copy())
        gm.max_radii2D = torch.from_numpy(s["max_radii2D"].copy())
        densify.densify_and_prune(gm, 0.0002, 0.005, 5.0, 20, None, noise=torch.from_numpy(noise))
        for g in dr.GROUPS:
            p = getattr(gm, ATTR[g])
            st = gm.optimizer.state[p]
            assert tuple(p.shape) == want[g].shape, g
            if g in ("xyz", "scaling"):
                np.testing.assert_allclose(p.detach().numpy(), want[g], rtol=2e-6, atol=2e-6)
            else:
                assert np.array_equal(p.detach().numpy(), want[g]), g
            assert np.array_equal(st["exp_avg"].numpy(), want["m_" + g])
            assert np.array_equal(st["exp_avg_sq"].numpy(), want["v_" + g])


@pytest.mark.skipif(not os.path.isdir("/root/reference/scene"), reason="needs the reference tree (authoring container)")
def test_install_on_the_reference_class_matches_its_own_methods(on_cpu, monkeypatch):
    """densify.install(GaussianModel) on the REFERENCE's class: the patched methods and the reference's originals run on
    two copies of the same model (CPU tensors; `plyfile` / `simple_knn` stubbed, device="cuda" redirected and
    torch.normal replaced by the stored table exactly as in tests/golden/make_densify_golden.py)."""
    import copy
    import importlib.util
    import sys
    spec = importlib.util.spec_from_file_location("make_densify_golden", os.path.join(HERE, "golden", "make_densify_golden.py"))
    saved = {n: getattr(torch, n) for n in ("zeros", "ones", "eye", "tensor", "empty", "normal")}
    saved_mods = {k: sys.modules.get(k) for k in ("plyfile", "simple_knn", "simple_knn._C", "scene", "scene.gaussian_model",
                                                  "utils", "utils.general_utils")}
    saved_path = list(sys.path)
    try:
        mk = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mk)                       # installs the stand-ins and imports the reference class
        gen = torch.Generator().manual_seed(77)
        ref = mk.make_model(300, gen)
        out = {}
        radii = mk.stats_steps(ref, gen, out, "t_")
        ours = copy.deepcopy(ref)
        Patched = type("Patched", (mk.GaussianModel,), {})
        densify.install(Patched)
        ours.__class__ = Patched
        Z = torch.randn(2, 600, 3, generator=gen)
        mk.NOISE["Z"], mk.NOISE["model"] = Z, ref
        ref.densify_and_prune(0.0002, 0.005, 5.0, 20, radii)
        ours.densify_and_prune(0.0002, 0.005, 5.0, 20, radii, noise=Z)
        ref.reset_opacity()
        ours.reset_opacity()
        mask = torch.rand(ref.get_xyz.shape[0], generator=gen) < 0.25
        ref.tmp_radii = torch.zeros(ref.get_xyz.shape[0])
        ref.prune_points(mask)
        ours.prune_points(mask)
        vsp = types.SimpleNamespace(grad=torch.randn(ref.get_xyz.shape[0], 3, generator=gen))
        flt = (torch.rand(ref.get_xyz.shape[0], generator=gen) < 0.5).nonzero()
        ref.add_densification_stats(vsp, flt)
        ours.add_densification_stats(vsp, flt)
        assert ours.get_xyz.shape == ref.get_xyz.shape and ref.get_xyz.shape[0] not in (0, 300)
        for a in ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation", "xyz_gradient_accum",
                  "denom", "max_radii2D"):
            assert torch.allclose(getattr(ours, a), getattr(ref, a), rtol=2e-6, atol=2e-6), a
        for go, gr in zip(ours.optimizer.param_groups, ref.optimizer.param_groups):
            so, sr = ours.optimizer.state[go["params"][0]], ref.optimizer.state[gr["params"][0]]
            assert go["name"] == gr["name"] and float(so["step"]) == float(sr["step"])
            assert torch.equal(so["exp_avg"], sr["exp_avg"]) and torch.equal(so["exp_avg_sq"], sr["exp_avg_sq"])
        # the activations of the reference class keep working on the rebound parameters
        assert ours.get_scaling.shape == ours._scaling.shape and ours.get_opacity.max() <= 0.0100001
        # and an optimizer step still updates the new parameter objects
        ours._xyz.grad = torch.ones_like(ours._xyz)
        before = ours._xyz.detach().clone()
        ours.optimizer.param_groups[0]["lr"] = 1e-3
        ours.optimizer.step()
        assert not torch.equal(ours._xyz.detach(), before)
    finally:
        for n, f in saved.items():
            setattr(torch, n, f)
        sys.path[:] = saved_path
        for k, m in saved_mods.items():
            if m is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = m
