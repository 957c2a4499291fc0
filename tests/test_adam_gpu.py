"""GPU tests of the SURVEY.md 8f rank-1 row: SparseGaussianAdam (fused kernel) and the `dc=` keyword
that render() uses once SparseGaussianAdam is importable (gaussian_renderer/__init__.py:115-125)."""
import numpy as np
import pytest
import torch

from tests.helpers import rel_err, settings_for, small_scene

pytestmark = pytest.mark.gpu


def _groups(P, dev, seed=0):
    g = torch.Generator().manual_seed(seed)
    shapes = {"xyz": (P, 3), "f_dc": (P, 1, 3), "f_rest": (P, 15, 3), "opacity": (P, 1), "scaling": (P, 3), "rotation": (P, 4)}
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 1.25e-4, "opacity": 0.025, "scaling": 0.005, "rotation": 0.001}
    params = {k: torch.nn.Parameter(torch.randn(*s, generator=g).to(dev)) for k, s in shapes.items()}
    return params, [{"params": [params[k]], "lr": lrs[k], "name": k} for k in shapes]


def test_sparse_adam_all_visible_equals_torch_adam():
    from diff_gaussian_rasterization import SparseGaussianAdam
    dev, P = torch.device("cuda"), 4097
    pa, ga = _groups(P, dev)
    pb, gb = _groups(P, dev)
    ours = SparseGaussianAdam(ga, lr=0.0, eps=1e-15)           # gaussian_model.py:248
    ref = torch.optim.Adam(gb, lr=0.0, eps=1e-15)              # gaussian_model.py:245
    vis = torch.ones(P, dtype=torch.bool, device=dev)
    g = torch.Generator().manual_seed(1)
    for step in range(5):
        for k in pa:
            grad = (torch.randn(pa[k].shape, generator=g) * 10.0 ** (step - 3)).to(dev)
            pa[k].grad, pb[k].grad = grad.clone(), grad.clone()
        ours.step(vis, P)
        ref.step()
    for k in pa:
        assert rel_err(pa[k].detach().cpu().numpy(), pb[k].detach().cpu().numpy()) < 1e-6, k
        sa, sb = ours.state[pa[k]], ref.state[pb[k]]
        assert rel_err(sa["exp_avg"].cpu().numpy(), sb["exp_avg"].cpu().numpy()) < 1e-6
        assert rel_err(sa["exp_avg_sq"].cpu().numpy(), sb["exp_avg_sq"].cpu().numpy()) < 1e-6
        assert float(sa["step"]) == float(sb["step"]) == 5.0


def test_sparse_adam_skips_invisible_rows():
    from diff_gaussian_rasterization import SparseGaussianAdam
    dev, P = torch.device("cuda"), 1000
    pa, ga = _groups(P, dev, seed=3)
    before = {k: v.detach().clone() for k, v in pa.items()}
    opt = SparseGaussianAdam(ga, lr=0.0, eps=1e-15)
    vis = torch.zeros(P, dtype=torch.bool, device=dev)
    vis[::3] = True
    for k in pa:
        pa[k].grad = torch.ones_like(pa[k])
    opt.step(vis, P)
    for k in pa:
        changed = (pa[k].detach() != before[k]).reshape(P, -1).any(dim=1)
        assert torch.equal(changed, vis), k
        m = opt.state[pa[k]]["exp_avg"].reshape(P, -1)
        assert torch.all(m[~vis] == 0) and torch.all(m[vis] != 0)


@pytest.mark.parametrize("P,deg,exchange", ((1500, 3, False), (4999, 3, False), (3001, 1, False), (2050, 2, False),
                                              (31, 3, False), (4999, 3, True), (2052, 1, True)))
def test_dc_keyword_equals_concatenated_shs(P, deg, exchange):
    """Split SH storage (`dc=` + `shs=`, gaussian_renderer/__init__.py:115-125) is read and differentiated in place by
    the kernels; same image bits and same gradients as the concatenated call.  Ragged Gaussian counts exercise the
    non-bulk tail warp; `exchange` runs the view-parallel factored exchange (one process) on the split layout."""
    import diff_gaussian_rasterization as dgr
    from gaussian_splatting_b200 import _lib, rasterizer
    W, H = 96, 64
    gs, s = small_scene(P=P, W=W, H=H, antialiasing=True, bg=(0.2, 0.1, 0.4), seed=17, sh_degree=deg,
                        scale_mult=6.0 if P > 100 else 30.0)
    dev = "cuda"
    rs = settings_for(dgr, s, dev)
    g = torch.Generator().manual_seed(2)
    dc_up = torch.randn(3, H, W, generator=g).to(dev) / 100
    dd_up = torch.randn(1, H, W, generator=g).to(dev) / 100

    def run(separate):
        t = {k: v.to(dev).requires_grad_(True) for k, v in gs.items() if k != "shs"}
        m2 = torch.zeros(P, 3, device=dev, requires_grad=True)
        f_dc = gs["shs"][:, :1, :].contiguous().to(dev).requires_grad_(True)
        f_rest = gs["shs"][:, 1:, :].contiguous().to(dev).requires_grad_(True)
        kw = dict(means3D=t["means3D"], means2D=m2, opacities=t["opacities"], scales=t["scales"], rotations=t["rotations"])
        r = dgr.GaussianRasterizer(rs)
        n0 = _lib.launch_count()
        color, radii, invd = r(dc=f_dc, shs=f_rest, **kw) if separate else r(shs=torch.cat((f_dc, f_rest), dim=1), **kw)
        torch.autograd.backward([color, invd], [dc_up, dd_up])
        out = dict(color=color.detach().cpu().numpy(), f_dc=f_dc.grad.cpu().numpy(), f_rest=f_rest.grad.cpu().numpy(),
                   m2=m2.grad.cpu().numpy())
        out.update({k: v.grad.cpu().numpy() for k, v in t.items()})
        return out, _lib.launch_count() - n0

    (b, _) = run(False)
    if exchange:
        from tests.test_exchange_gpu import LocalExchange
        rasterizer.set_gradient_reducer(LocalExchange())
    try:
        (a, launches) = run(True)
    finally:
        rasterizer.set_gradient_reducer(None)
    assert np.array_equal(a["color"], b["color"])
    for k in b:
        if k != "color":
            assert rel_err(a[k], b[k]) < 1e-5, k
    ncoef = (deg + 1) ** 2
    assert a["f_dc"].shape == (P, 1, 3) and a["f_rest"].shape == (P, ncoef - 1, 3)
    assert launches > 10          # our kernels ran (no torch.cat path: nothing but libgsrast launches is needed)
