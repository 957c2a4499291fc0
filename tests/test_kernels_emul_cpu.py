"""The product's CUDA kernels executed on the CPU (no GPU in the authoring container) against the C oracle.

tests/host_emul/cuda_emul provides a stand-in `cuda_runtime.h` and a fiber scheduler (every CUDA thread of a block is a
fiber; __syncthreads, __shfl_*_sync, __ballot_sync, ... are scheduling points; bulk async copies complete at issue).
ALL of gaussian_splatting_b200/csrc/*.cu is compiled as C++ against it into tests/host_emul/libgsrast_emul.so, which
exports the same C-ABI as libgsrast.so; this test drives it the way a non-Python host would (plain pointers, numpy
buffers): gsr_forward_stage1 -> R -> gsr_forward_stage2 -> gsr_backward, gsr_mark_visible, gsr_adam_update, gsr_ssim_*.
Checked against oracle/ with the tolerances of the GPU parity tests: radii / num_rendered / point_list / ranges /
n_contrib exact, images 1e-5, gradients 1e-3 relative.  What the emulation cannot show (races, asynchrony, speed) is
the business of the -m gpu tests; what it does show is that the kernel sources compute the reference's results."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
import torch

from gaussian_splatting_b200 import _lib
from gaussian_splatting_b200._lib import GsrBackwardIO, GsrBuffers, GsrForwardOut, GsrParams
from tests.helpers import np_inputs, np_settings, rel_err, run_oracle, small_scene

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "..", "gaussian_splatting_b200", "csrc")
EMU = os.path.join(HERE, "host_emul", "cuda_emul")
UNITS = ("api", "preprocess", "binning", "render_fwd", "render_bwd", "preprocess_bwd", "adam", "ssim", "exchange", "densify", "peak")
VP = C.c_void_p


def load_full_emul(defines=(), tag="full"):
    """defines / tag: a variant of the emulated library (e.g. a lowered threshold) next to the default one."""
    out = os.path.join(HERE, "host_emul", f"libgsrast_{tag}_emul.so")
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(EMU, f) for f in ("cuda_runtime.h", "emul.cpp", "async_copy_emul.cuh")]
    deps.append(os.path.join(HERE, "..", "include", "gsrast.h"))
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(d) for d in deps):
        objs = []
        procs = []
        for u in UNITS:
            o = os.path.join(HERE, "host_emul", f"_emul_{tag}_{u}.o")
            objs.append(o)
            procs.append(subprocess.Popen(["g++", "-std=c++17", "-O1", "-fPIC", "-ffp-contract=off", "-U_FORTIFY_SOURCE", "-D_FORTIFY_SOURCE=0", "-DGSR_HOST_EMUL",
                                           *[f"-D{d}" for d in defines], "-I" + EMU, "-c", "-x", "c++", os.path.join(CSRC, u + ".cu"), "-o", o]))
        assert all(p.wait() == 0 for p in procs)
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-shared", "-fPIC", "-U_FORTIFY_SOURCE", "-D_FORTIFY_SOURCE=0", "-I" + EMU,
                               os.path.join(EMU, "emul.cpp")]
                              + objs + ["-o", out])
    lib = C.CDLL(out)
    PP, PB, PO = C.POINTER(GsrParams), C.POINTER(GsrBuffers), C.POINTER(GsrForwardOut)
    lib.gsr_last_error.restype = C.c_char_p
    for n, a in (("gsr_geom_bytes", [C.c_int32]), ("gsr_img_bytes", [C.c_int32, C.c_int32]), ("gsr_bin_bytes", [C.c_int64]),
                 ("gsr_scratch_bytes", [C.c_int32, C.c_int32, C.c_int32, C.c_int64])):
        getattr(lib, n).restype = C.c_size_t
        getattr(lib, n).argtypes = a
    lib.gsr_forward_stage1.argtypes = [PP, PB, PO, VP]
    lib.gsr_forward_stage2.argtypes = [PP, PB, PO, C.c_int64, VP]
    lib.gsr_backward.argtypes = [PP, PB, C.c_int64, VP, C.POINTER(GsrBackwardIO), VP]
    lib.gsr_mark_visible.argtypes = [C.c_int32, VP, VP, VP, VP, VP]
    lib.gsr_debug_state_ptrs.argtypes = [PP, PB, C.c_int64] + [C.POINTER(VP)] * 5
    lib.gsr_debug_sorted_keys.argtypes = [PP, PB, C.c_int64, VP, VP]
    lib.gsr_adam_update.argtypes = [VP] * 5 + [C.c_double] * 6 + [C.c_int64, C.c_int32, VP]
    PS = C.POINTER(_lib.GsrSsimParams)
    lib.gsr_ssim_partials.restype = C.c_int64
    lib.gsr_ssim_partials.argtypes = [C.c_int32] * 4
    lib.gsr_ssim_forward.argtypes = [PS] + [VP] * 7
    lib.gsr_ssim_backward.argtypes = [PS] + [VP] * 7
    return lib


@pytest.fixture(scope="module")
def klib():
    return load_full_emul()


def ok(lib, rc):
    assert rc == 0, lib.gsr_last_error().decode()


def aligned(nbytes, align=256):
    # state and scratch arrive uninitialised on the device (torch.empty): fill with garbage, not zeros
    raw = np.full(int(nbytes) + align, 0xA5, np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + int(nbytes)]


def f32(a):
    return np.ascontiguousarray(np.asarray(a, np.float32))


class EmulRasterizer:
    """One forward (+ backward) through the emulated library, buffers owned here as a C host would own them."""

    def __init__(self, lib, gs, s, split_sh=False):
        self.lib = lib
        inp, st = np_inputs(gs), np_settings(s)
        self.P = P = inp["means3D"].shape[0]
        self.W, self.H = int(st["image_width"]), int(st["image_height"])
        self.keep = k = {n: f32(inp[n]) for n in inp}
        k["bg"], k["view"], k["proj"], k["campos"] = f32(st["bg"]), f32(st["viewmatrix"]), f32(st["projmatrix"]), f32(st["campos"])
        p = self.p = GsrParams()
        p.P, p.D, p.W, p.H = P, int(st["sh_degree"]), self.W, self.H
        p.tanfovx, p.tanfovy, p.scale_modifier = float(st["tanfovx"]), float(st["tanfovy"]), float(st["scale_modifier"])
        p.prefiltered, p.antialiasing, p.debug = 0, int(bool(st["antialiasing"])), 0
        ptr = lambda a: a.ctypes.data
        p.background, p.means3D, p.opacities = ptr(k["bg"]), ptr(k["means3D"]), ptr(k["opacities"])
        p.viewmatrix, p.projmatrix, p.campos = ptr(k["view"]), ptr(k["proj"]), ptr(k["campos"])
        self.split = split_sh
        if "shs" in k:
            p.M = k["shs"].shape[1]
            if split_sh:
                k["dc"], k["rest"] = f32(k["shs"][:, :1]), f32(k["shs"][:, 1:])
                p.sh_dc, p.shs = ptr(k["dc"]), ptr(k["rest"])
            else:
                p.shs = ptr(k["shs"])
        else:
            p.colors_precomp = ptr(k["colors_precomp"])
        if "scales" in k:
            p.scales, p.rotations = ptr(k["scales"]), ptr(k["rotations"])
        else:
            p.cov3D_precomp = ptr(k["cov3D_precomp"])
        b = self.b = GsrBuffers()
        self.geom, self.img = aligned(lib.gsr_geom_bytes(P)), aligned(lib.gsr_img_bytes(self.W, self.H))
        self.scratch = aligned(lib.gsr_scratch_bytes(P, self.W, self.H, 0))
        self.pinned = np.zeros(2, np.int32)
        b.geom, b.geom_bytes, b.img, b.img_bytes = ptr(self.geom), self.geom.nbytes, ptr(self.img), self.img.nbytes
        b.scratch, b.scratch_bytes, b.num_rendered_host = ptr(self.scratch), self.scratch.nbytes, ptr(self.pinned)
        self.color, self.invdepth = np.zeros((3, self.H, self.W), np.float32), np.zeros((1, self.H, self.W), np.float32)
        self.radii = np.zeros(P, np.int32)
        self.o = GsrForwardOut(ptr(self.color), ptr(self.invdepth), ptr(self.radii))

    def forward(self):
        lib, p, b, o = self.lib, self.p, self.b, self.o
        ok(lib, lib.gsr_forward_stage1(C.byref(p), C.byref(b), C.byref(o), None))
        self.R = R = int(self.pinned[0])
        assert self.pinned[1] == 0
        need = lib.gsr_scratch_bytes(self.P, self.W, self.H, R)
        if need > self.scratch.nbytes:                 # regrow keeps the stage-1 prefix
            new = aligned(need)
            new[:self.scratch.nbytes] = self.scratch
            self.scratch = new
            b.scratch, b.scratch_bytes = new.ctypes.data, new.nbytes
        self.bin = aligned(lib.gsr_bin_bytes(R))
        b.bin, b.bin_bytes = self.bin.ctypes.data, self.bin.nbytes
        ok(lib, lib.gsr_forward_stage2(C.byref(p), C.byref(b), C.byref(o), R, None))
        return self

    def state(self):
        ptrs = [VP() for _ in range(5)]
        ok(self.lib, self.lib.gsr_debug_state_ptrs(C.byref(self.p), C.byref(self.b), self.R, *[C.byref(x) for x in ptrs]))
        T = ((self.W + 15) // 16) * ((self.H + 15) // 16)
        N = self.W * self.H
        arr = lambda q, n, dt: np.ctypeslib.as_array(C.cast(q, C.POINTER(C.c_uint32)), (n,)).view(dt).copy()
        keys = np.zeros(max(self.R, 1), np.uint64)
        ok(self.lib, self.lib.gsr_debug_sorted_keys(C.byref(self.p), C.byref(self.b), self.R, keys.ctypes.data, None))
        return dict(point_list=arr(ptrs[0], self.R, np.uint32), ranges=arr(ptrs[1], 2 * T, np.uint32).reshape(T, 2),
                    final_T=arr(ptrs[2], N, np.float32), n_contrib=arr(ptrs[3], N, np.uint32), keys=keys[:self.R])

    def backward(self, dc, dd):
        P, M = self.P, self.p.M
        g = {"means3D": np.empty((P, 3), np.float32), "means2D": np.empty((P, 3), np.float32),
             "opacities": np.empty((P, 1), np.float32)}
        io = GsrBackwardIO()
        self.dc, self.dd = f32(dc), (None if dd is None else f32(dd))
        io.dL_dout_color = self.dc.ctypes.data
        io.dL_dout_invdepth = None if dd is None else self.dd.ctypes.data
        io.dL_dmeans3D, io.dL_dmeans2D, io.dL_dopacity = (g[n].ctypes.data for n in ("means3D", "means2D", "opacities"))
        if self.p.shs:
            if self.split:
                g["dc"], g["rest"] = np.empty((P, 1, 3), np.float32), np.empty((P, M - 1, 3), np.float32)
                io.dL_dsh_dc, io.dL_dsh = g["dc"].ctypes.data, g["rest"].ctypes.data
            else:
                g["shs"] = np.empty((P, M, 3), np.float32)
                io.dL_dsh = g["shs"].ctypes.data
        else:
            g["colors_precomp"] = np.empty((P, 3), np.float32)
            io.dL_dcolors = g["colors_precomp"].ctypes.data
        if self.p.scales:
            g["scales"], g["rotations"] = np.empty((P, 3), np.float32), np.empty((P, 4), np.float32)
            io.dL_dscales, io.dL_drotations = g["scales"].ctypes.data, g["rotations"].ctypes.data
        else:
            g["cov3D_precomp"] = np.empty((P, 6), np.float32)
            io.dL_dcov3D = g["cov3D_precomp"].ctypes.data
        for a in g.values():
            a.fill(np.nan)                       # every output must be fully written (gsrast.h)
        ok(self.lib, self.lib.gsr_backward(C.byref(self.p), C.byref(self.b), self.R, self.radii.ctypes.data, C.byref(io), None))
        if self.split:
            g["shs"] = np.concatenate((g.pop("dc"), g.pop("rest")), 1)
        return g


def upstream(W, H, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(3, H, W, generator=g) / 100, torch.randn(1, H, W, generator=g) / 100


def check_against_oracle(lib, oracle, gs, s, dd_on=True, split_sh=False, state=True):
    W, H = int(s["image_width"]), int(s["image_height"])
    dc, dd = upstream(W, H)
    want = run_oracle(oracle, gs, s, dc, dd if dd_on else None)
    er = EmulRasterizer(lib, gs, s, split_sh).forward()
    assert np.array_equal(er.radii, want["radii"])
    fw = want["fw"]
    assert er.R == int(fw["R"])
    assert np.abs(er.color - want["color"]).max() <= 1e-5 and np.abs(er.invdepth - want["invdepth"]).max() <= 1e-5
    if state:
        st = er.state()
        assert np.array_equal(st["point_list"], fw["point_list"][:er.R])
        assert np.array_equal(st["ranges"], np.asarray(fw["ranges"]).reshape(-1, 2))
        assert np.array_equal(st["n_contrib"], np.asarray(fw["n_contrib"]).reshape(-1))
        assert np.abs(st["final_T"] - np.asarray(fw["final_T"]).reshape(-1)).max() <= 1e-5
        assert np.array_equal(st["keys"], fw["keys"])          # tile << 32 | depth bits, sorted
    got = er.backward(dc.numpy(), dd.numpy() if dd_on else None)
    for k, g in got.items():
        assert np.isfinite(g).all(), k
        assert rel_err(g, np.asarray(want["grads"][k]).reshape(g.shape)) < 1e-3, k
    return er


def test_forward_backward_sh_scale_rotation(klib, oracle):
    gs, s = small_scene(P=600, W=64, H=48, seed=1, antialiasing=True, bg=(0.1, 0.3, 0.6))
    check_against_oracle(klib, oracle, gs, s)


def test_split_sh_storage_and_no_depth_gradient(klib, oracle):
    gs, s = small_scene(P=300, W=48, H=32, seed=2, sh_degree=2)
    check_against_oracle(klib, oracle, gs, s, dd_on=False, split_sh=True, state=False)


def test_precomputed_colors_and_covariances(klib, oracle):
    gs, s = small_scene(P=300, W=40, H=40, seed=3, sh_degree=0, bg=(1.0, 1.0, 1.0))
    g = torch.Generator().manual_seed(9)
    gs = dict(gs)
    gs["colors_precomp"] = torch.rand(300, 3, generator=g)
    del gs["shs"]
    # a valid covariance: the one the scales / rotations describe (computeCov3D layout xx xy xz yy yz zz)
    q, sc = gs.pop("rotations"), gs.pop("scales")
    w, x, y, z = q.unbind(1)
    R = torch.stack((1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                     2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                     2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)), 1).reshape(-1, 3, 3)
    Sg = R @ torch.diag_embed(sc * sc) @ R.transpose(1, 2)
    gs["cov3D_precomp"] = torch.stack((Sg[:, 0, 0], Sg[:, 0, 1], Sg[:, 0, 2], Sg[:, 1, 1], Sg[:, 1, 2], Sg[:, 2, 2]), 1).contiguous()
    check_against_oracle(klib, oracle, gs, s, state=False)


def test_ragged_image_and_many_batches(klib, oracle):
    # image not a multiple of the tile size; > 256 entries in a tile => several double-buffered batches per tile
    gs, s = small_scene(P=1500, W=37, H=21, seed=4, scale_mult=10.0, behind_fraction=0.0)
    er = check_against_oracle(klib, oracle, gs, s)
    assert er.R / 6 > 256


def test_mark_visible(klib, oracle):
    gs, s = small_scene(P=500, W=32, H=32, seed=5, behind_fraction=0.4)
    means, view, proj = f32(gs["means3D"]), f32(s["viewmatrix"]), f32(s["projmatrix"])
    present = np.zeros(500, np.uint8)
    ok(klib, klib.gsr_mark_visible(500, means.ctypes.data, view.ctypes.data, proj.ctypes.data, present.ctypes.data, None))
    z = means @ view[:3, 2] + view[3, 2]                       # (W2V p).z with the column-major matrix
    assert np.array_equal(present.astype(bool), z > 0.2) and 0 < present.sum() < 500


def test_adam_kernel_matches_torch(klib):
    g = torch.Generator().manual_seed(0)
    N, M = 257, 45
    p = torch.randn(N, M, generator=g)
    grad, m, v = torch.randn(N, M, generator=g), torch.zeros(N, M), torch.zeros(N, M)
    ref = torch.nn.Parameter(p.clone())
    opt = torch.optim.Adam([ref], lr=1e-2, eps=1e-15)
    vis = (torch.rand(N, generator=g) < 0.6).to(torch.uint8)
    pa, ga, ma, va = (f32(t.numpy()) for t in (p, grad, m, v))
    for step in (1, 2, 3):
        ref.grad = grad * vis[:, None]
        opt.step()
        ok(klib, klib.gsr_adam_update(pa.ctypes.data, ga.ctypes.data, ma.ctypes.data, va.ctypes.data,
                                      vis.numpy().ctypes.data, 1e-2, 0.9, 0.999, 1e-15, 1 - 0.9 ** step,
                                      (1 - 0.999 ** step) ** 0.5, N, M, None))
    visb = vis.bool().numpy()
    # visible rows: torch.optim.Adam's update; invisible rows untouched (torch's Adam moves them: zero gradient
    # still decays the moments, which is the documented difference of the sparse optimizer)
    assert rel_err(pa[visb], ref.detach().numpy()[visb]) < 1e-6
    assert np.array_equal(pa[~visb], p.numpy()[~visb]) and not ma[~visb].any()


def test_ssim_kernels_match_the_oracle(klib):
    from oracle import ssim_ref
    rng = np.random.default_rng(0)
    B, CH, H, W = 1, 2, 45, 70
    a, b = rng.random((B, CH, H, W), dtype=np.float32), rng.random((B, CH, H, W), dtype=np.float32)
    sp = _lib.GsrSsimParams(a.ctypes.data, b.ctypes.data, B, CH, H, W, 0.01 ** 2, 0.03 ** 2, 0)
    maps = [np.zeros_like(a) for _ in range(4)]
    mean = np.zeros(1, np.float32)
    partials = np.zeros(int(klib.gsr_ssim_partials(B, CH, H, W)), np.float32)
    ok(klib, klib.gsr_ssim_forward(C.byref(sp), *[m.ctypes.data for m in maps], mean.ctypes.data, partials.ctypes.data, None))
    v_ref, g_ref = ssim_ref.fused_ssim(a, b, dtype=np.float64)
    m_ref, *_ = ssim_ref.ssim_forward(a, b, dtype=np.float64)
    assert abs(float(mean[0]) - v_ref) < 1e-6 and np.abs(maps[0] - m_ref).max() < 1e-5
    dmean = np.ones(1, np.float32)
    dimg = np.zeros_like(a)
    ok(klib, klib.gsr_ssim_backward(C.byref(sp), None, dmean.ctypes.data, maps[1].ctypes.data, maps[2].ctypes.data,
                                    maps[3].ctypes.data, dimg.ctypes.data, None))
    assert rel_err(dimg, g_ref) < 1e-4


@pytest.mark.parametrize("deg,split", ((3, False), (3, True), (2, False), (1, True), (0, False)))
def test_factored_exchange_two_ranks(klib, oracle, deg, split):
    """View-parallel exchange with world = 2 on the CPU: two cameras of the same Gaussians run the sender
    branch of preprocess_bwd.cu with multimem_reduce = 1 — under the emulation a multimem.red is an add into the one
    local buffer, i.e. exactly the sum the NVSwitch forms in every replica — then gsr_exchange_expand rebuilds the dense
    gradients; they must equal the sum of the two cameras' ordinary gradients (oracle)."""
    P, W, H = 400, 48, 32
    klib.gsr_exchange_expand.argtypes = [C.POINTER(GsrParams), C.c_int32] + [VP] * 10
    world = 2
    records = np.zeros((P, 12), np.float32)
    rgb = np.zeros((world, P, 4), np.float32)
    campos = np.zeros((world, 4), np.float32)
    want = None
    dc, dd = upstream(W, H, seed=3)
    for r, yaw in enumerate((0.0, 25.0)):
        gs, s = small_scene(P=P, W=W, H=H, seed=6, yaw_deg=yaw, antialiasing=True, behind_fraction=0.0, sh_degree=deg)   # replicated Gaussians
        o = run_oracle(oracle, gs, s, dc, dd)
        want = o["grads"] if want is None else {k: want[k] + o["grads"][k] for k in want if k != "means2D"}
        er = EmulRasterizer(klib, gs, s, split_sh=split).forward()
        io = GsrBackwardIO()
        dcn, ddn = f32(dc.numpy()), f32(dd.numpy())
        m2 = np.full((P, 3), np.nan, np.float32)
        io.dL_dout_color, io.dL_dout_invdepth, io.dL_dmeans2D = dcn.ctypes.data, ddn.ctypes.data, m2.ctypes.data
        io.multimem_reduce = 1
        io.xchg_records, io.xchg_rgb, io.xchg_campos = records.ctypes.data, rgb[r].ctypes.data, campos[r].ctypes.data
        ok(klib, klib.gsr_backward(C.byref(er.p), C.byref(er.b), er.R, er.radii.ctypes.data, C.byref(io), None))
        assert rel_err(m2, o["grads"]["means2D"]) < 1e-3       # per-view statistic, never reduced
        assert np.allclose(campos[r], np.append(f32(s["campos"]), 1.0))
    out = {"means3D": np.full((P, 3), np.nan, np.float32), "opacities": np.full((P, 1), np.nan, np.float32),
           "shs": np.full((P, (deg + 1) ** 2, 3), np.nan, np.float32), "scales": np.full((P, 3), np.nan, np.float32),
           "rotations": np.full((P, 4), np.nan, np.float32)}
    sh_dc = np.full((P, 1, 3), np.nan, np.float32)
    sh_rest = np.full((P, (deg + 1) ** 2 - 1, 3), np.nan, np.float32)
    ok(klib, klib.gsr_exchange_expand(C.byref(er.p), world, records.ctypes.data, rgb.ctypes.data, campos.ctypes.data,
                                      out["means3D"].ctypes.data, out["opacities"].ctypes.data,
                                      sh_rest.ctypes.data if split else out["shs"].ctypes.data,
                                      out["scales"].ctypes.data, out["rotations"].ctypes.data,
                                      sh_dc.ctypes.data if split else None, None))
    if split:                                    # split SH storage: coefficient 0 and the rest leave separately
        out["shs"] = np.concatenate((sh_dc, sh_rest), 1)
    for k, g in out.items():
        assert rel_err(g, np.asarray(want[k]).reshape(g.shape)) < 1e-3, k
    # self-cleaning buffers: the rows expand consumed are zero again for the next step
    assert not records.any() and not rgb.any()


@pytest.mark.parametrize("bands", (0, 3))
def test_full_multicast_reduction_with_band_pipelining(klib, oracle, bands):
    """multimem_reduce = 1 without the factored exchange (`GSR_REDUCE_FACTORED=0`): every gradient store is
    an add into the (pre-zeroed) replica; with `reserved1 = bands` the blend backward runs tile-row band by band and
    each band's Gaussians go to the per-Gaussian kernel through index lists.  One rank here, so the result must be the
    ordinary gradient."""
    P, W, H = 500, 48, 64
    gs, s = small_scene(P=P, W=W, H=H, seed=8, antialiasing=True)
    dc, dd = upstream(W, H, seed=5)
    want = run_oracle(oracle, gs, s, dc, dd)["grads"]
    er = EmulRasterizer(klib, gs, s).forward()
    g = {"means3D": np.zeros((P, 3), np.float32), "opacities": np.zeros((P, 1), np.float32),
         "shs": np.zeros((P, 16, 3), np.float32), "scales": np.zeros((P, 3), np.float32),
         "rotations": np.zeros((P, 4), np.float32)}
    m2 = np.full((P, 3), np.nan, np.float32)
    io = GsrBackwardIO()
    dcn, ddn = f32(dc.numpy()), f32(dd.numpy())
    io.dL_dout_color, io.dL_dout_invdepth, io.dL_dmeans2D = dcn.ctypes.data, ddn.ctypes.data, m2.ctypes.data
    io.dL_dmeans3D, io.dL_dopacity, io.dL_dsh = g["means3D"].ctypes.data, g["opacities"].ctypes.data, g["shs"].ctypes.data
    io.dL_dscales, io.dL_drotations = g["scales"].ctypes.data, g["rotations"].ctypes.data
    io.multimem_reduce, io.reserved1 = 1, bands
    ok(klib, klib.gsr_backward(C.byref(er.p), C.byref(er.b), er.R, er.radii.ctypes.data, C.byref(io), None))
    assert rel_err(m2, want["means2D"]) < 1e-3
    for k, v in g.items():
        assert rel_err(v, np.asarray(want[k]).reshape(v.shape)) < 1e-3, k


def test_debug_mode_and_error_reporting(klib):
    gs, s = small_scene(P=100, W=32, H=32, seed=9)
    er = EmulRasterizer(klib, gs, s)
    er.p.debug = 1                                   # synchronise + check after every stage
    er.forward()
    assert er.R > 0
    bad = EmulRasterizer(klib, gs, s)
    bad.p.shs = None                                 # neither SHs nor precomputed colours
    rc = klib.gsr_forward_stage1(C.byref(bad.p), C.byref(bad.b), C.byref(bad.o), None)
    assert rc == _lib.GSR_ERR_INVALID_ARG
    assert klib.gsr_last_error().decode() == "Please provide exactly one of either SHs or precomputed colors!"
    small = EmulRasterizer(klib, gs, s)
    small.b.geom_bytes = 64
    assert klib.gsr_forward_stage1(C.byref(small.p), C.byref(small.b), C.byref(small.o), None) == _lib.GSR_ERR_CAPACITY


def test_more_than_65536_tiles_nine_bands(klib, oracle):
    """257 x 257 tiles: more tiles than one unit of the tile-list construction keeps in shared memory, so the image is
    cut into 9 bands of 31 tile rows (binning.cu) — the path the GPU suite covers with test_large_image_many_bands.
    Small Gaussians keep the emulation at about a minute."""
    gs, s = small_scene(P=300, W=4112, H=4112, seed=5, scale_mult=0.1, antialiasing=True)
    er = check_against_oracle(klib, oracle, gs, s)
    assert ((er.W + 15) // 16) * ((er.H + 15) // 16) > 65536 and er.R > 10000


def test_radix_fallback_pass_for_counts_beyond_30_bits(oracle):
    """Above 2^30 items the onesweep look-back counters (30 bits) do not suffice and binning.cu falls back to a
    histogram / row-scan / scatter pass per digit with full 32-bit counters — unreachable at test sizes, so this variant
    of the emulated library lowers the threshold (-DGSR_OS_MAX_N=64) and both sorts take the fallback."""
    lib = load_full_emul(defines=("GSR_OS_MAX_N=64",), tag="fallback")
    gs, s = small_scene(P=700, W=80, H=48, seed=13, scale_mult=4.0, antialiasing=True)
    er = check_against_oracle(lib, oracle, gs, s)
    assert er.P > 64 and er.R > 64


@pytest.mark.parametrize("P,W,H,scale", [(700, 80, 48, 4.0), (2500, 200, 120, 1.0), (130, 33, 17, 30.0)])
def test_tile_lists_many_chunks_and_bands(oracle, P, W, H, scale):
    """The tile-list construction at small sizes with its shared-memory budget lowered to 16 tiles (-DGSR_TB_MAX_CELLS=16):
    bands of one or two tile rows, several chunks per band, rectangles that straddle bands and (scale 30) cover the whole
    image with more than 32 tiles per Gaussian.  point_list, ranges and the rebuilt keys stay bit-exact."""
    lib = load_full_emul(defines=("GSR_TB_MAX_CELLS=16",), tag="bands")
    gs, s = small_scene(P=P, W=W, H=H, seed=21, scale_mult=scale, antialiasing=True)
    er = check_against_oracle(lib, oracle, gs, s)
    assert er.R > 64


def test_tile_lists_screen_filling_gaussians(klib, oracle):
    """Screen-filling Gaussians: every rectangle reaches every stripe of the list-filling kernel and the clipped
    rectangles exceed 32 tiles (several passes per step, binning.cu)."""
    gs, s = small_scene(P=900, W=200, H=136, seed=27, scale_mult=25.0, antialiasing=True)
    er = check_against_oracle(klib, oracle, gs, s)
    assert er.R > 900 * 40


def test_adam_scalar_path_and_unmasked(klib):
    # element count not a multiple of 4 -> the scalar kernel instead of the 16-byte one; visible == NULL -> everywhere
    g = torch.Generator().manual_seed(1)
    N, M = 7, 3
    p = torch.randn(N, M, generator=g)
    grad = torch.randn(N, M, generator=g)
    ref = torch.nn.Parameter(p.clone())
    opt = torch.optim.Adam([ref], lr=3e-3, eps=1e-15)
    pa, ga, ma, va = f32(p.numpy()), f32(grad.numpy()), np.zeros((N, M), np.float32), np.zeros((N, M), np.float32)
    for step in (1, 2):
        ref.grad = grad.clone()
        opt.step()
        ok(klib, klib.gsr_adam_update(pa.ctypes.data, ga.ctypes.data, ma.ctypes.data, va.ctypes.data, None, 3e-3, 0.9, 0.999,
                                      1e-15, 1 - 0.9 ** step, (1 - 0.999 ** step) ** 0.5, N, M, None))
    assert rel_err(pa, ref.detach().numpy()) < 1e-6
    assert rel_err(ma, opt.state[ref]["exp_avg"].numpy()) < 1e-6


@pytest.mark.parametrize("shape", ((2, 1, 23, 40), (1, 3, 12, 11)))
def test_ssim_valid_crop_and_map_gradient(klib, shape):
    """padding="valid" (crop = 5) through the fused-mean path, and the FusedSSIMMap path (map gradient in)."""
    from oracle import ssim_ref
    rng = np.random.default_rng(2)
    B, CH, H, W = shape
    a, b = rng.random(shape, dtype=np.float32), rng.random(shape, dtype=np.float32)
    sp = _lib.GsrSsimParams(a.ctypes.data, b.ctypes.data, B, CH, H, W, 0.01 ** 2, 0.03 ** 2, 5)
    maps = [np.zeros_like(a) for _ in range(4)]
    mean = np.zeros(1, np.float32)
    partials = np.zeros(int(klib.gsr_ssim_partials(B, CH, H, W)), np.float32)
    ok(klib, klib.gsr_ssim_forward(C.byref(sp), *[m.ctypes.data for m in maps], mean.ctypes.data, partials.ctypes.data, None))
    v_ref, g_ref = ssim_ref.fused_ssim(a, b, "valid", dtype=np.float64)
    assert abs(float(mean[0]) - v_ref) < 2e-6
    dmean = np.ones(1, np.float32)
    dimg = np.full_like(a, np.nan)
    ok(klib, klib.gsr_ssim_backward(C.byref(sp), None, dmean.ctypes.data, maps[1].ctypes.data, maps[2].ctypes.data,
                                    maps[3].ctypes.data, dimg.ctypes.data, None))
    assert rel_err(dimg, g_ref) < 1e-4
    # map gradient in (crop 0): a random dL_dmap must give sum(dL_dmap * dmap/dimg1), linear in dL_dmap
    sp0 = _lib.GsrSsimParams(a.ctypes.data, b.ctypes.data, B, CH, H, W, 0.01 ** 2, 0.03 ** 2, 0)
    ok(klib, klib.gsr_ssim_forward(C.byref(sp0), *[m.ctypes.data for m in maps], None, None, None))
    dmap = rng.random(shape, dtype=np.float32)
    d1, d2 = np.zeros_like(a), np.zeros_like(a)
    ok(klib, klib.gsr_ssim_backward(C.byref(sp0), dmap.ctypes.data, None, maps[1].ctypes.data, maps[2].ctypes.data,
                                    maps[3].ctypes.data, d1.ctypes.data, None))
    dmap2 = f32(2 * dmap)
    ok(klib, klib.gsr_ssim_backward(C.byref(sp0), dmap2.ctypes.data, None, maps[1].ctypes.data, maps[2].ctypes.data,
                                    maps[3].ctypes.data, d2.ctypes.data, None))
    assert rel_err(d2, 2 * d1) < 1e-6
    ones = np.full(shape, 1.0 / a.size, np.float32)
    ok(klib, klib.gsr_ssim_backward(C.byref(sp0), ones.ctypes.data, None, maps[1].ctypes.data, maps[2].ctypes.data,
                                    maps[3].ctypes.data, d1.ctypes.data, None))
    _, g_same = ssim_ref.fused_ssim(a, b, "same", dtype=np.float64)
    assert rel_err(d1, g_same) < 1e-4


def test_one_call_forward_capacity_protocol_and_work_counters(klib, oracle):
    """gsr_forward (stage 1, wait, stage 2 in one call): with a bin buffer that is too small it reports
    GSR_ERR_CAPACITY and still returns num_rendered, the caller regrows and finishes with gsr_forward_stage2;
    gsr_debug_count_work (the bench's work descriptors Wf / Bp / Wb) against the oracle's own counters."""
    klib.gsr_forward.argtypes = [C.POINTER(GsrParams), C.POINTER(GsrBuffers), C.POINTER(GsrForwardOut),
                                 C.POINTER(C.c_int64), VP]
    klib.gsr_debug_count_work.argtypes = [C.POINTER(GsrParams), C.POINTER(GsrBuffers), C.c_int64, VP, VP]
    gs, s = small_scene(P=400, W=48, H=48, seed=14, antialiasing=True)
    want = run_oracle(oracle, gs, s)
    er = EmulRasterizer(klib, gs, s)
    er.scratch = aligned(klib.gsr_scratch_bytes(er.P, er.W, er.H, 1 << 16))       # generous: only bin is short
    er.b.scratch, er.b.scratch_bytes = er.scratch.ctypes.data, er.scratch.nbytes
    tiny = aligned(256)
    er.b.bin, er.b.bin_bytes = tiny.ctypes.data, tiny.nbytes
    R = C.c_int64(-1)
    rc = klib.gsr_forward(C.byref(er.p), C.byref(er.b), C.byref(er.o), C.byref(R), None)
    assert rc == _lib.GSR_ERR_CAPACITY and R.value == int(want["fw"]["R"]) > 64
    er.R = R.value
    er.bin = aligned(klib.gsr_bin_bytes(er.R))
    er.b.bin, er.b.bin_bytes = er.bin.ctypes.data, er.bin.nbytes
    ok(klib, klib.gsr_forward_stage2(C.byref(er.p), C.byref(er.b), C.byref(er.o), er.R, None))
    assert np.abs(er.color - want["color"]).max() <= 1e-5
    # the same call with enough room succeeds in one go
    R2 = C.c_int64(-1)
    ok(klib, klib.gsr_forward(C.byref(er.p), C.byref(er.b), C.byref(er.o), C.byref(R2), None))
    assert R2.value == er.R and np.abs(er.color - want["color"]).max() <= 1e-5
    counts = np.zeros(3, np.uint64)
    ok(klib, klib.gsr_debug_count_work(C.byref(er.p), C.byref(er.b), er.R, counts.ctypes.data, None))
    stats = want["fw"]["stats"]
    assert int(counts[0]) == int(stats[0]) and int(counts[1]) == int(stats[1])
    assert int(counts[2]) == int(np.asarray(want["fw"]["n_contrib"], np.int64).sum())


@pytest.mark.parametrize("deg,P,split", ((0, 97, False), (1, 64, False), (2, 130, False), (3, 301, True), (1, 35, True)))
def test_sh_layouts_of_the_per_gaussian_backward(klib, oracle, deg, P, split):
    """M = 1, 4, 9, 16 coefficients: rows that are / are not whole float4s take the vector / scalar staging of
    preprocess_bwd.cu; ragged last warps of the split layout are copied by the lanes instead of bulk copies."""
    gs, s = small_scene(P=P, W=40, H=24, seed=20 + deg, sh_degree=deg, antialiasing=bool(deg & 1))
    check_against_oracle(klib, oracle, gs, s, split_sh=split, state=False)


@pytest.mark.parametrize("masked", (True, False))
def test_adam_vector_kernel(klib, masked):
    # element count a multiple of 4 and 16-byte aligned tensors: the float4 kernel every real model tensor takes
    g = torch.Generator().manual_seed(5)
    N, M = 256, 45
    p, grad = torch.randn(N, M, generator=g), torch.randn(N, M, generator=g)
    vis = (torch.rand(N, generator=g) < 0.5).to(torch.uint8) if masked else torch.ones(N, dtype=torch.uint8)
    ref = torch.nn.Parameter(p.clone())
    opt = torch.optim.Adam([ref], lr=1e-2, eps=1e-15)
    bufs = [aligned(N * M * 4, 16).view(np.float32).reshape(N, M) for _ in range(4)]
    pa, ga, ma, va = bufs
    pa[:], ga[:], ma[:], va[:] = p.numpy(), grad.numpy(), 0, 0
    for step in (1, 2, 3):
        ref.grad = grad * vis[:, None]
        opt.step()
        ok(klib, klib.gsr_adam_update(pa.ctypes.data, ga.ctypes.data, ma.ctypes.data, va.ctypes.data,
                                      vis.numpy().ctypes.data if masked else None, 1e-2, 0.9, 0.999, 1e-15,
                                      1 - 0.9 ** step, (1 - 0.999 ** step) ** 0.5, N, M, None))
    v = vis.bool().numpy()
    assert rel_err(pa[v], ref.detach().numpy()[v]) < 1e-6
    assert np.array_equal(pa[~v], p.numpy()[~v]) and not ma[~v].any() and not va[~v].any()
    assert klib.gsr_adam_update(None, None, None, None, None, 1e-2, 0.9, 0.999, 1e-15, 0.1, 0.1, 4, 3, None) == _lib.GSR_ERR_INVALID_ARG


def test_profiling_hooks_and_argument_errors(klib):
    klib.gsr_profile_collect.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    klib.gsr_stage_name.restype = C.c_char_p
    klib.gsr_launch_count.restype = C.c_uint64
    gs, s = small_scene(P=120, W=32, H=32, seed=15)
    klib.gsr_profile_enable(1)
    l0 = klib.gsr_launch_count()
    er = EmulRasterizer(klib, gs, s).forward()
    er.backward(np.ones((3, 32, 32), np.float32), None)
    ms, n = (C.c_double * _lib.STAGE_COUNT)(), (C.c_int64 * _lib.STAGE_COUNT)()
    ok(klib, klib.gsr_profile_collect(ms, n))
    klib.gsr_profile_enable(0)
    names = [klib.gsr_stage_name(i).decode() for i in range(_lib.STAGE_COUNT)]
    assert all(c == 1 for c, name in zip(n, names) if name != "exchange") and klib.gsr_stage_name(99) == b"?"
    assert names == ["preprocess", "depth_sort", "tile_count", "tile_fill", "render_fwd", "render_bwd", "preprocess_bwd",
                     "exchange"]
    assert klib.gsr_launch_count() - l0 >= 13
    # argument errors come back as codes with the reference's messages, never as aborts
    e = EmulRasterizer(klib, gs, s)
    e.p.scales = None
    assert klib.gsr_forward_stage1(C.byref(e.p), C.byref(e.b), C.byref(e.o), None) == _lib.GSR_ERR_INVALID_ARG
    assert b"scale/rotation pair or precomputed 3D covariance" in klib.gsr_last_error()
    e = EmulRasterizer(klib, gs, s)
    e.p.M = 3                                            # fewer coefficients than the active degree needs
    assert klib.gsr_forward_stage1(C.byref(e.p), C.byref(e.b), C.byref(e.o), None) == _lib.GSR_ERR_INVALID_ARG
    e = EmulRasterizer(klib, gs, s)
    e.b.img_bytes = 16
    assert klib.gsr_forward_stage1(C.byref(e.p), C.byref(e.b), C.byref(e.o), None) == _lib.GSR_ERR_CAPACITY
    e = EmulRasterizer(klib, gs, s)
    e.o.out_color = None
    rc1 = klib.gsr_forward_stage1(C.byref(e.p), C.byref(e.b), C.byref(e.o), None)
    assert rc1 == _lib.GSR_ERR_INVALID_ARG or klib.gsr_forward_stage2(C.byref(e.p), C.byref(e.b), C.byref(e.o), 0, None) != 0


def test_ssim_inference_mode_and_argument_errors(klib):
    from oracle import ssim_ref
    rng = np.random.default_rng(4)
    a, b = rng.random((1, 1, 20, 33), dtype=np.float32), rng.random((1, 1, 20, 33), dtype=np.float32)
    sp = _lib.GsrSsimParams(a.ctypes.data, b.ctypes.data, 1, 1, 20, 33, 0.01 ** 2, 0.03 ** 2, 0)
    smap = np.zeros_like(a)
    ok(klib, klib.gsr_ssim_forward(C.byref(sp), smap.ctypes.data, None, None, None, None, None, None))   # train=False
    m_ref, *_ = ssim_ref.ssim_forward(a, b, dtype=np.float64)
    assert np.abs(smap - m_ref).max() < 1e-5
    one = np.zeros_like(a)
    assert klib.gsr_ssim_forward(C.byref(sp), smap.ctypes.data, one.ctypes.data, None, None, None, None, None) == _lib.GSR_ERR_INVALID_ARG
    mean = np.zeros(1, np.float32)
    assert klib.gsr_ssim_forward(C.byref(sp), None, None, None, None, mean.ctypes.data, None, None) == _lib.GSR_ERR_INVALID_ARG
    assert klib.gsr_ssim_backward(C.byref(sp), smap.ctypes.data, None, None, None, None, one.ctypes.data, None) == _lib.GSR_ERR_INVALID_ARG
    bad = _lib.GsrSsimParams(None, b.ctypes.data, 1, 1, 20, 33, 1e-4, 9e-4, 0)
    assert klib.gsr_ssim_forward(C.byref(bad), smap.ctypes.data, None, None, None, None, None, None) == _lib.GSR_ERR_INVALID_ARG
