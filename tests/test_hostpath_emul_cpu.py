"""The product's PYTHON host layers (autograd Function, workspace handling, SparseGaussianAdam, fused_ssim, density
control, PLY I/O) executed in the GPU-less container on top of the emulated kernel library.

`_lib` is pointed at tests/host_emul/libgsrast_full_emul.so (all csrc/*.cu compiled against the CUDA emulation, see
tests/test_kernels_emul_cpu.py) and — in this test process only — the handful of torch.cuda entry points the host code
touches are replaced by inert stand-ins (current_stream / device / Event / pin_memory / Tensor.is_cuda), so the SAME
Python code that runs on the B200 runs here on CPU tensors.  The product itself has no CPU path: without these
test-only patches it refuses CPU tensors (asserted below)."""
import contextlib
import types

import numpy as np
import pytest
import torch

from oracle import ssim_ref
from tests.helpers import rel_err, run_module, run_oracle, small_scene
from tests.test_kernels_emul_cpu import load_full_emul


class _Stream:
    cuda_stream = 0

    def synchronize(self):
        pass

    def wait_event(self, ev):
        pass

    def wait_stream(self, s):
        pass

    def record_event(self, ev=None):
        return ev or _Event()


class _Event:
    cuda_event = None        # the emulated library records nothing (its stream is synchronous)

    def __init__(self, *a, **k):
        pass

    def record(self, stream=None):
        pass

    def synchronize(self):
        pass

    def wait(self, stream=None):
        pass

    def elapsed_time(self, other):
        return 0.0

    def query(self):
        return True


@pytest.fixture()
def emulated_cuda(monkeypatch):
    import os
    from gaussian_splatting_b200 import _lib, densify, plyio, rasterizer
    load_full_emul()                                   # build if stale
    here = os.path.dirname(os.path.abspath(__file__))
    monkeypatch.setattr(_lib, "LIB_PATH", os.path.join(here, "host_emul", "libgsrast_full_emul.so"))
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(torch.cuda, "current_stream", lambda device=None: _Stream())
    monkeypatch.setattr(torch.cuda, "device", lambda d: contextlib.nullcontext())
    monkeypatch.setattr(torch.cuda, "current_device", lambda: 0)
    monkeypatch.setattr(torch.cuda, "synchronize", lambda device=None: None)
    monkeypatch.setattr(torch.cuda, "Event", _Event)
    monkeypatch.setattr(torch.Tensor, "pin_memory", lambda self, *a, **k: self)
    monkeypatch.setattr(torch.Tensor, "is_cuda", property(lambda self: True))
    monkeypatch.setattr(rasterizer, "_workspaces", {})
    for mod in (densify, plyio):
        monkeypatch.setattr(mod, "_device_ctx", lambda d: contextlib.nullcontext())
    yield
    _lib._lib = None


def test_the_product_refuses_cpu_tensors_without_the_patches():
    import diff_gaussian_rasterization as dgr
    from tests.helpers import settings_for
    gs, s = small_scene(P=10, W=16, H=16)
    rast = dgr.GaussianRasterizer(settings_for(dgr, s, "cpu"))
    with pytest.raises(RuntimeError, match="CUDA"):
        rast(means3D=gs["means3D"], means2D=torch.zeros(10, 3), opacities=gs["opacities"], shs=gs["shs"],
             scales=gs["scales"], rotations=gs["rotations"])


@pytest.mark.parametrize("aa,bg", ((True, (0.1, 0.3, 0.6)), (False, (0.0, 0.0, 0.0))))
def test_autograd_path_matches_oracle(emulated_cuda, oracle, aa, bg):
    import diff_gaussian_rasterization as dgr
    W, H = 64, 48
    gs, s = small_scene(P=500, W=W, H=H, seed=11, antialiasing=aa, bg=bg)
    g = torch.Generator().manual_seed(0)
    dc, dd = torch.randn(3, H, W, generator=g) / 100, torch.randn(1, H, W, generator=g) / 100
    got = run_module(dgr, gs, s, dc, dd, device="cpu")
    want = run_oracle(oracle, gs, s, dc, dd)
    assert np.array_equal(got["radii"], want["radii"])
    assert np.abs(got["color"] - want["color"]).max() <= 1e-5 and np.abs(got["invdepth"] - want["invdepth"]).max() <= 1e-5
    for k in ("means3D", "means2D", "opacities", "shs", "scales", "rotations"):
        assert rel_err(got["grads"][k], want["grads"][k].reshape(got["grads"][k].shape)) < 1e-3, k
    # a second call reuses the workspace arena; P == 0 returns zero images without a launch
    again = run_module(dgr, gs, s, dc, dd, device="cpu")
    assert np.array_equal(again["color"], got["color"])
    empty = {k: v[:0] for k, v in gs.items()}
    z = run_module(dgr, empty, s, device="cpu")
    assert not z["color"].any() and z["radii"].shape == (0,)


def test_dc_keyword_equals_concatenated_call(emulated_cuda):
    import diff_gaussian_rasterization as dgr
    from tests.helpers import settings_for
    W, H = 48, 32
    gs, s = small_scene(P=301, W=W, H=H, seed=12, sh_degree=2)          # ragged: 301 % 4 != 0
    rast = dgr.GaussianRasterizer(settings_for(dgr, s, "cpu"))
    outs = []
    for split in (False, True):
        leaves = {k: v.clone().requires_grad_(True) for k, v in gs.items()}
        dcl = gs["shs"][:, :1].clone().requires_grad_(True)
        rest = gs["shs"][:, 1:].clone().requires_grad_(True)
        m2 = torch.zeros(301, 3, requires_grad=True)
        kw = dict(dc=dcl, shs=rest) if split else dict(shs=leaves["shs"])
        color, radii, invd = rast(means3D=leaves["means3D"], means2D=m2, opacities=leaves["opacities"],
                                  scales=leaves["scales"], rotations=leaves["rotations"], **kw)
        (color.sum() + invd.sum()).backward()
        gsh = torch.cat((dcl.grad, rest.grad), 1) if split else leaves["shs"].grad
        outs.append((color.detach(), gsh, leaves["means3D"].grad, leaves["scales"].grad))
    assert torch.equal(outs[0][0], outs[1][0])
    for a, b in zip(outs[0][1:], outs[1][1:]):
        assert rel_err(a.numpy(), b.numpy()) < 1e-5


def _raw_model_params(gs, seed=0):
    """The model's raw leaves for a scene's activated values: log-scales, quaternions of arbitrary length, logits."""
    g = torch.Generator().manual_seed(seed)
    raw_s = gs["scales"].log()
    raw_r = gs["rotations"] * (0.3 + 2.0 * torch.rand(gs["rotations"].shape[0], 1, generator=g))
    o = gs["opacities"].clamp(1e-4, 1 - 1e-4)
    raw_o = (o / (1 - o)).log()
    return raw_s, raw_r, raw_o


@pytest.mark.parametrize("how", ["keyword", "deferred"])
def test_fused_activations_match_torch_autograd(emulated_cuda, how):
    """GsrParams.raw_inputs: the kernels activate the raw parameters themselves and return raw gradients.  Oracle: the
    reference's own activations (torch.exp / normalize / sigmoid, scene/gaussian_model.py:44-55) + torch autograd around
    the unfused call."""
    import diff_gaussian_rasterization as dgr
    from gaussian_splatting_b200 import activations as act
    from tests.helpers import settings_for
    W, H, P = 64, 48, 400
    gs, s = small_scene(P=P, W=W, H=H, seed=17, antialiasing=True, bg=(0.2, 0.1, 0.4))
    raw_s, raw_r, raw_o = _raw_model_params(gs)
    rast = dgr.GaussianRasterizer(settings_for(dgr, s, "cpu"))
    g = torch.Generator().manual_seed(1)
    dc, dd = torch.randn(3, H, W, generator=g) / 100, torch.randn(1, H, W, generator=g) / 100
    outs = []
    for fused in (False, True):
        ls, lr, lo = (t.clone().requires_grad_(True) for t in (raw_s, raw_r, raw_o))
        lm, lsh = gs["means3D"].clone().requires_grad_(True), gs["shs"].clone().requires_grad_(True)
        m2 = torch.zeros(P, 3, requires_grad=True)
        if not fused:
            kw = dict(opacities=torch.sigmoid(lo), scales=torch.exp(ls), rotations=torch.nn.functional.normalize(lr))
        elif how == "keyword":
            kw = dict(opacities=lo, scales=ls, rotations=lr, raw_inputs=act.RAW_SCALES | act.RAW_ROTATIONS | act.RAW_OPACITIES)
        else:
            kw = dict(opacities=act.Deferred(lo, torch.sigmoid, act.RAW_OPACITIES), scales=act.Deferred(ls, torch.exp, act.RAW_SCALES),
                      rotations=act.Deferred(lr, torch.nn.functional.normalize, act.RAW_ROTATIONS))
        color, radii, invd = rast(means3D=lm, means2D=m2, shs=lsh, **kw)
        torch.autograd.backward([color, invd], [dc, dd])
        outs.append(dict(color=color.detach(), invd=invd.detach(), radii=radii, s=ls.grad, r=lr.grad, o=lo.grad,
                         m=lm.grad, sh=lsh.grad, m2=m2.grad))
    a, b = outs
    assert torch.equal(a["radii"], b["radii"])
    assert (a["color"] - b["color"]).abs().max() <= 1e-5 and (a["invd"] - b["invd"]).abs().max() <= 1e-5
    for k in ("s", "r", "o", "m", "sh", "m2"):
        assert rel_err(b[k].numpy(), a[k].numpy()) < 1e-4, k


def test_deferred_activation_behaves_like_the_tensor_elsewhere():
    """What the reference does with get_scaling / get_opacity outside render() (gaussian_model.py:324,521-535,612-619)
    works on the proxies that activations.install() makes the properties return; a model with other activation
    functions keeps the reference's behaviour."""
    from gaussian_splatting_b200 import activations as act

    class Model:                                       # the three properties of scene/gaussian_model.py:130-157
        def __init__(self):
            g = torch.Generator().manual_seed(0)
            self._scaling = torch.randn(50, 3, generator=g)
            self._rotation = torch.randn(50, 4, generator=g)
            self._opacity = torch.randn(50, 1, generator=g)
            self.scaling_activation, self.opacity_activation = torch.exp, torch.sigmoid
            self.rotation_activation = torch.nn.functional.normalize

        @property
        def get_scaling(self):
            return self.scaling_activation(self._scaling)

        @property
        def get_rotation(self):
            return self.rotation_activation(self._rotation)

        @property
        def get_opacity(self):
            return self.opacity_activation(self._opacity)

    ref = Model()
    want = dict(s=ref.get_scaling, r=ref.get_rotation, o=ref.get_opacity)
    act.install(Model)
    try:
        m = Model()
        s, r, o = m.get_scaling, m.get_rotation, m.get_opacity
        assert all(isinstance(t, act.Deferred) for t in (s, r, o))
        assert torch.equal(torch.max(s, dim=1).values, torch.max(want["s"], dim=1).values)
        mask = torch.max(s, dim=1).values > 1.0
        assert torch.equal(s[mask].repeat(2, 1), want["s"][mask].repeat(2, 1))
        assert torch.equal(s.max(dim=1).values > 0.5, want["s"].max(dim=1).values > 0.5)
        assert torch.equal((o < 0.4).squeeze(), (want["o"] < 0.4).squeeze())
        assert torch.equal(torch.min(o, torch.ones_like(o) * 0.01), torch.min(want["o"], torch.ones_like(want["o"]) * 0.01))
        assert torch.equal(s / 1.6, want["s"] / 1.6) and o.shape == (50, 1) and len(r) == 50
        assert torch.equal(r.value(), want["r"])
        o2, s2, r2, bits = act.unwrap(o, s, r)
        assert o2 is m._opacity and s2 is m._scaling and r2 is m._rotation and bits == 7
        m.scaling_activation = torch.nn.functional.softplus          # not what the kernels implement: no proxy
        assert isinstance(m.get_scaling, torch.Tensor) and isinstance(m.get_opacity, torch.Tensor)
    finally:
        act.uninstall(Model)
    assert isinstance(Model().get_scaling, torch.Tensor)


def test_sparse_adam_and_fused_ssim(emulated_cuda):
    import diff_gaussian_rasterization as dgr
    import fused_ssim as fs
    g = torch.Generator().manual_seed(3)
    pa = torch.nn.Parameter(torch.randn(130, 45, generator=g))
    pb = torch.nn.Parameter(pa.detach().clone())
    ours = dgr.SparseGaussianAdam([{"params": [pa], "lr": 1e-2, "name": "f_rest"}], lr=0.0, eps=1e-15)
    ref = torch.optim.Adam([{"params": [pb], "lr": 1e-2}], lr=0.0, eps=1e-15)
    for _ in range(3):
        grad = torch.randn(130, 45, generator=g)
        pa.grad, pb.grad = grad.clone(), grad.clone()
        ours.step(torch.ones(130, dtype=torch.bool), 130)
        ref.step()
    assert rel_err(pa.detach().numpy(), pb.detach().numpy()) < 1e-6
    img1 = torch.rand(1, 3, 40, 52, generator=g).requires_grad_(True)
    img2 = torch.rand(1, 3, 40, 52, generator=g)
    v = fs.fused_ssim(img1, img2)
    v.backward()
    v_ref, g_ref = ssim_ref.fused_ssim(img1.detach().numpy(), img2.numpy(), dtype=np.float64)
    assert abs(float(v.detach()) - v_ref) < 1e-5 and rel_err(img1.grad.numpy(), g_ref) < 1e-3


@pytest.mark.parametrize("depth,mask", [(False, False), (True, True)])
def test_fused_training_loss_matches_torch_and_the_ssim_oracle(emulated_cuda, depth, mask):
    """loss.training_loss (one forward + one backward launch) on the emulation: L1 and inverse-depth terms against
    torch, the SSIM term against oracle/ssim_ref.py (fp64), gradients against torch autograd + the oracle's gradient."""
    from gaussian_splatting_b200 import loss as L
    H, W, lam, wd = 37, 52, 0.2, 0.37
    g = torch.Generator().manual_seed(5)
    image = torch.rand(3, H, W, generator=g)
    gt = (image + 0.2 * torch.randn(3, H, W, generator=g)).clamp(0, 1)
    gt[:, :10] = image[:, :10]
    invd = torch.rand(1, H, W, generator=g) if depth else None
    mono = torch.rand(1, H, W, generator=g) if depth else None
    msk = (torch.rand(1, H, W, generator=g) > 0.3).float() if mask else None
    a_img = image.clone().requires_grad_(True)
    a_inv = invd.clone().requires_grad_(True) if depth else None
    loss, parts = L.training_loss(a_img, gt, lam, invdepth=a_inv, mono_invdepth=mono, depth_mask=msk, depth_weight=wd,
                                  return_parts=True)
    (loss * 1.7).backward()
    v_ref, g_ref = ssim_ref.fused_ssim(image[None].numpy(), gt[None].numpy(), dtype=np.float64)
    b_img = image.clone().requires_grad_(True)
    b_inv = invd.clone().requires_grad_(True) if depth else None
    l1 = (b_img - gt).abs().mean()
    want = (1 - lam) * l1 + lam * (1 - float(v_ref))
    d = torch.zeros(())
    if depth:
        d = ((b_inv - mono) * msk).abs().mean()
        want = want + wd * d
    (want * 1.7).backward()
    grad_want = b_img.grad.numpy() - 1.7 * lam * g_ref[0]
    assert abs(float(loss) - float(want)) < 2e-6 and abs(float(parts[0]) - float(l1)) < 2e-6
    assert abs(float(parts[1]) - float(v_ref)) < 2e-6 and abs(float(parts[2]) - float(d)) < 2e-6
    assert rel_err(a_img.grad.numpy(), grad_want) < 1e-4
    if depth:
        assert rel_err(a_inv.grad.numpy(), b_inv.grad.numpy()) < 1e-6


def test_training_slice_of_the_gpu_suite(emulated_cuda, tmp_path, monkeypatch):
    """tests/test_zz_densify_gpu.py::test_training_slice_render_stats_step_densify_save_load — the callers either side
    of the rasterizer in one piece — executed here on the emulation."""
    from gaussian_splatting_b200 import densify, plyio
    from tests import test_zz_densify_gpu as G
    monkeypatch.setattr(G, "DEV", "cpu")
    monkeypatch.setattr(densify, "_pinned_counts", lambda: torch.zeros(4, dtype=torch.int32))
    monkeypatch.setattr(plyio, "_to_host", lambda rows: rows)
    monkeypatch.setattr(plyio, "_to_device", lambda host, device: host)
    G.test_training_slice_render_stats_step_densify_save_load(tmp_path)
