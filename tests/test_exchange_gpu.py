"""Single-GPU tests of the factored gradient exchange of the view-parallel step:
sender branch of preprocess_bwd.cu (xchg_* buffers) + receiver csrc/exchange.cu.

The multicast adds themselves need >= 2 GPUs (tests/multi_gpu_check.py, run by test_parity_gpu's
multi-GPU wrapper); here the same kernels run with plain stores and the cross-rank sum of the 48 B
records is emulated with a torch add, which exercises everything else: packing, the per-rank colour
slots, the camera slots and the reconstruction  dL_dsh = sum_r basis(mean - campos_r) x dL_dRGB_r.
Oracle: the ordinary backward of the same views (itself checked against the reference CUDA build).
"""
import ctypes as C

import numpy as np
import pytest
import torch

from tests.helpers import rel_err, run_module, small_scene

pytestmark = pytest.mark.gpu


class LocalExchange:
    """Test double of view_parallel.FusedGradientReducer for ONE process: same interface, ordinary device
    memory, plain stores (multicast = False)."""
    multicast = False
    world = 1
    bands = 1

    def __init__(self):
        self.flat = None
        self.begun = self.ended = self.prepared = 0

    def use_exchange(self, names):
        return [n for n in names if n != "sh_dc"] == ["means3D", "opacity", "sh", "scales", "rotations"]

    def exchange_buffers(self, P, device):
        n = 12 * P + 4 * P + 4
        if self.flat is None or self.flat.numel() != n:
            self.flat = torch.empty(n, device=device)
        b = self.flat.data_ptr()
        d = dict(records=b, rgb=b + 48 * P, campos=b + 64 * P)
        return d, d

    def prepare(self, shapes, device, exchange_P=None):
        assert exchange_P is not None
        self.exchange_buffers(exchange_P, device)
        self.prepared += 1

    def begin(self):
        self.flat.zero_()
        self.begun += 1

    def end(self):
        self.ended += 1


@pytest.mark.parametrize("deg,aa", ((3, True), (1, False), (0, False)))
def test_exchange_world1_equals_plain_backward(deg, aa):
    import diff_gaussian_rasterization as dgr
    from gaussian_splatting_b200 import rasterizer
    W, H = 112, 80
    gs, s = small_scene(P=3000, W=W, H=H, antialiasing=aa, bg=(0.2, 0.1, 0.4), seed=3, sh_degree=deg)
    g = torch.Generator().manual_seed(0)
    dc, dd = torch.randn(3, H, W, generator=g) / 100, torch.randn(1, H, W, generator=g) / 100
    want = run_module(dgr, gs, s, dc, dd)
    red = LocalExchange()
    rasterizer.set_gradient_reducer(red)
    try:
        got = run_module(dgr, gs, s, dc, dd)
    finally:
        rasterizer.set_gradient_reducer(None)
    assert (red.prepared, red.begun, red.ended) == (1, 1, 1)
    assert np.array_equal(got["color"], want["color"])
    for k, v in want["grads"].items():
        # the blend backward accumulates with atomics, so two runs differ in the last bits
        assert rel_err(got["grads"][k], v) < 1e-5, k


def test_exchange_two_views_emulated():
    """Two cameras over the same Gaussians: records added, one colour / camera slot per view."""
    import diff_gaussian_rasterization as dgr
    from gaussian_splatting_b200 import _lib, rasterizer
    from tests.helpers import settings_for
    lib = _lib.load()
    W, H, P = 96, 64, 2500
    views = [small_scene(P=P, W=W, H=H, antialiasing=True, seed=5, yaw_deg=y, behind_fraction=0.0) for y in (0.0, 70.0)]
    gs = views[0][0]
    g = torch.Generator().manual_seed(1)
    dc, dd = torch.randn(3, H, W, generator=g) / 100, torch.randn(1, H, W, generator=g) / 100
    want = None
    for _, s in views:
        r = run_module(dgr, gs, s, dc, dd)["grads"]
        want = r if want is None else {k: want[k] + r[k] for k in r}

    dev = torch.device("cuda")
    leaves = {k: v.to(dev) for k, v in gs.items()}
    records = torch.zeros(P, 12, device=dev)
    rgb = torch.zeros(2, P, 4, device=dev)
    campos = torch.zeros(2, 4, device=dev)
    e = torch.Tensor([])
    for v, (_, s) in enumerate(views):
        rs = settings_for(dgr, s, dev)
        R, color, radii, geom, binb, img, invd = rasterizer._forward_native(
            rs, leaves["means3D"], leaves["shs"], e, leaves["opacities"], leaves["scales"], leaves["rotations"], e)
        rec_v = torch.zeros(P, 12, device=dev)
        dummy = torch.zeros(P * 16 + 4, device=dev)     # what this "rank" expands locally: nothing

        class OneView(LocalExchange):
            def exchange_buffers(self, P_, device):
                # the kernel publishes into this view's slots ("multicast" side); the rasterizer's own expand
                # call reads the dummy replica, so the published data survives for the emulated reduction below
                pub = dict(records=rec_v.data_ptr(), rgb=rgb[v].data_ptr(), campos=campos[v].data_ptr())
                b = dummy.data_ptr()
                return dict(records=b, rgb=b + 48 * P_, campos=b + 64 * P_), pub

            def begin(self):
                self.begun += 1
        red = OneView()
        rasterizer.set_gradient_reducer(red)
        try:
            rasterizer._backward_native(rs, leaves["means3D"], radii, e, leaves["opacities"], leaves["scales"],
                                        leaves["rotations"], e, dc.to(dev), dd.to(dev), leaves["shs"], geom, R, binb, img)
        finally:
            rasterizer.set_gradient_reducer(None)
        records += rec_v                                  # what multimem.red.add does across ranks
    assert torch.allclose(campos[:, :3].cpu(), torch.stack([s["campos"] for _, s in views])) and (campos[:, 3] == 1).all()
    # a Gaussian seen by both views really gets two contributions
    both = ((rgb[0, :, :3].abs().sum(1) > 0) & (rgb[1, :, :3].abs().sum(1) > 0)).sum().item()
    assert both > 50
    outs = {k: torch.empty_like(leaves[k]) for k in ("means3D", "opacities", "shs", "scales", "rotations")}
    p = _lib.GsrParams()
    p.P, p.D, p.M, p.means3D = P, 3, 16, leaves["means3D"].data_ptr()
    _lib.check(lib.gsr_exchange_expand(C.byref(p), 2, records.data_ptr(), rgb.data_ptr(), campos.data_ptr(),
                                       outs["means3D"].data_ptr(), outs["opacities"].data_ptr(), outs["shs"].data_ptr(),
                                       outs["scales"].data_ptr(), outs["rotations"].data_ptr(), None,
                                       torch.cuda.current_stream().cuda_stream))
    for k, t in outs.items():
        assert rel_err(t.cpu().numpy(), want[k].reshape(t.shape)) < 1e-5, k
    # the expand pass leaves the exchange buffers clean for the next step (no per-step memset)
    assert not records.any() and not rgb.any()


def test_exchange_argument_checks():
    from gaussian_splatting_b200 import _lib
    lib = _lib.load()
    p = _lib.GsrParams()
    p.P, p.D, p.M = 10, 3, 16
    buf = torch.zeros(1024, device="cuda")
    a = buf.data_ptr()
    assert lib.gsr_exchange_expand(C.byref(p), 0, a, a, a, a, a, None, a, a, None, None) == _lib.GSR_ERR_INVALID_ARG
    assert lib.gsr_exchange_expand(C.byref(p), 2, a + 4, a, a, a, a, None, a, a, None, None) == _lib.GSR_ERR_INVALID_ARG
    assert b"aligned" in lib.gsr_last_error()
    p.D, p.M = 3, 4
    assert lib.gsr_exchange_expand(C.byref(p), 2, a, a, a, a, a, None, a, a, None, None) == _lib.GSR_ERR_INVALID_ARG
    p.P = 0
    p.M = 16
    assert lib.gsr_exchange_expand(C.byref(p), 1, None, None, None, None, None, None, None, None, None, None) == _lib.GSR_OK
