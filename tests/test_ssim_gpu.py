"""GPU parity tests of the fused SSIM kernels (csrc/ssim.cu through the C-ABI) — SURVEY.md 8f rank 2.

Checked against (a) the numpy oracle (oracle/ssim_ref.py, fp64) on the golden inputs and on ragged
shapes, (b) the golden values from the reference's Python SSIM, (c) the reference's own ssim.cu built
for sm_100 (oracle/_ref/fused_ssim), (d) the conv2d formulation with torch.isclose exactly as the
reference's tests/test.py:57-91 does, plus size-independent properties at full size.
Tolerances (floating point): map 1e-5 absolute (north_star's image tolerance), gradients 1e-3 relative
to the largest gradient entry; measured errors are ~1e-7 / ~1e-5.
"""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import ref_loader, ssim_ref
from tests.helpers import rel_err

pytestmark = pytest.mark.gpu

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ssim_golden.npz"))
C1, C2 = 0.01 ** 2, 0.03 ** 2
MAP_TOL, GRAD_TOL = 1e-5, 1e-3


def _pair(shape, seed, dev="cuda"):
    g = torch.Generator().manual_seed(seed)
    a = torch.rand(shape, generator=g)
    b = (0.7 * a + 0.3 * torch.rand(shape, generator=g)).clamp(0, 1)
    return a.to(dev), b.to(dev)


def _conv_ssim(img1, img2):
    """utils/loss_utils.py:56-86 (what tests/test.py calls the reference implementation)."""
    ch = img1.size(-3)
    g = torch.from_numpy(ssim_ref.taps(np.float32)).to(img1.device)
    w = (g[:, None] @ g[None, :]).expand(ch, 1, 11, 11).contiguous()
    mu1, mu2 = F.conv2d(img1, w, padding=5, groups=ch), F.conv2d(img2, w, padding=5, groups=ch)
    s1 = F.conv2d(img1 * img1, w, padding=5, groups=ch) - mu1.pow(2)
    s2 = F.conv2d(img2 * img2, w, padding=5, groups=ch) - mu2.pow(2)
    s12 = F.conv2d(img1 * img2, w, padding=5, groups=ch) - mu1 * mu2
    return (((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1.pow(2) + mu2.pow(2) + C1) * (s1 + s2 + C2))).mean()


@pytest.mark.parametrize("case", ("a", "b", "c", "d"))
def test_map_and_derivative_maps_vs_oracle(case):
    import fused_ssim as fs
    from gaussian_splatting_b200 import ssim as impl
    a, b = GOLD[f"{case}_img1"], GOLD[f"{case}_img2"]
    ta, tb = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    smap, _, dms, _, _ = impl._forward(C1, C2, ta, tb, 0, True, True, False)
    want = ssim_ref.ssim_forward(a, b, dtype=np.float64)
    assert np.abs(smap.cpu().numpy() - want[0]).max() < MAP_TOL
    assert np.abs(smap.cpu().numpy() - GOLD[f"{case}_map_f64"]).max() < MAP_TOL      # the reference's own numbers
    for got, ref in zip(dms, want[1:]):
        assert rel_err(got.cpu().numpy(), ref) < 1e-4
    # autograd surface, map variant with a random upstream gradient
    g = torch.Generator().manual_seed(1)
    dl = torch.randn(a.shape, generator=g).cuda()
    x = ta.clone().requires_grad_(True)
    m = fs.FusedSSIMMap.apply(C1, C2, x, tb, "same", True)
    assert torch.equal(m, smap)
    m.backward(dl)
    ref = ssim_ref.ssim_backward(a, b, dl.cpu().numpy(), *want[1:], dtype=np.float64)
    assert rel_err(x.grad.cpu().numpy(), ref) < GRAD_TOL


@pytest.mark.parametrize("case", ("a", "b", "c", "d"))
@pytest.mark.parametrize("padding", ("same", "valid"))
def test_fused_mean_vs_reference_python_golden(case, padding):
    import fused_ssim as fs
    a, b = GOLD[f"{case}_img1"], GOLD[f"{case}_img2"]
    if padding == "valid" and (a.shape[2] <= 10 or a.shape[3] <= 10):
        pytest.skip("empty crop")
    x = torch.from_numpy(a).cuda().requires_grad_(True)
    v = fs.fused_ssim(x, torch.from_numpy(b).cuda(), padding)
    (3.0 * v).backward()                                  # non-unit upstream scalar
    assert abs(float(v.detach()) - float(GOLD[f"{case}_{padding}_value_f64"])) < MAP_TOL
    assert rel_err(x.grad.cpu().numpy() / 3.0, GOLD[f"{case}_{padding}_grad_f64"]) < GRAD_TOL
    # the map variant (FusedSSIMMap + crop + mean in torch, as the reference wrapper) agrees with the fused mean
    x2 = torch.from_numpy(a).cuda().requires_grad_(True)
    v2 = fs.FusedSSIMMap.apply(C1, C2, x2, torch.from_numpy(b).cuda(), padding, True).mean()
    v2.backward()
    assert abs(float(v2.detach()) - float(v.detach())) < 1e-6
    assert rel_err(x2.grad.cpu().numpy(), x.grad.cpu().numpy() / 3.0) < 1e-5


@pytest.mark.parametrize("shape", ((1, 1, 1, 1), (1, 2, 3, 40), (2, 1, 33, 2), (1, 3, 31, 65), (3, 4, 64, 32), (1, 3, 97, 130)))
def test_ragged_shapes_vs_oracle(shape):
    import fused_ssim as fs
    a, b = _pair(shape, seed=sum(shape))
    x = a.clone().requires_grad_(True)
    v = fs.fused_ssim(x, b)
    v.backward()
    v_ref, g_ref = ssim_ref.fused_ssim(a.cpu().numpy(), b.cpu().numpy(), dtype=np.float64)
    assert abs(float(v) - float(v_ref)) < MAP_TOL
    assert rel_err(x.grad.cpu().numpy(), g_ref) < GRAD_TOL


def test_matches_conv2d_formulation_like_the_reference_test():
    """tests/test.py:57-91: torch.isclose on the value and on every gradient element."""
    import fused_ssim as fs
    torch.manual_seed(0)
    for shape in ((2, 5, 270, 480), (1, 3, 545, 980)):
        img1 = torch.nn.Parameter(torch.rand(shape, device="cuda"))
        img2 = torch.rand(shape, device="cuda")
        mine = torch.nn.Parameter(img1.detach().clone())
        og = _conv_ssim(img1, img2)
        val = fs.fused_ssim(mine, img2)
        assert torch.isclose(og, val)
        og.backward()
        val.backward()
        assert torch.isclose(img1.grad, mine.grad).all()


@pytest.mark.skipif(not ref_loader.fused_ssim_available(), reason="oracle/_ref/fused_ssim not built")
@pytest.mark.parametrize("shape,padding", (((1, 3, 1080, 1920), "same"), ((2, 5, 360, 641), "valid"), ((1, 3, 840, 1297), "same")))
def test_vs_reference_cuda_build(shape, padding):
    """The reference's own ssim.cu (compiled for sm_100) on the same inputs, at the benchmark sizes."""
    import fused_ssim as fs
    ref = ref_loader.load_fused_ssim()
    a, b = _pair(shape, seed=7)
    x_ref = a.clone().requires_grad_(True)
    v_ref = ref.fused_ssim(x_ref, b, padding)
    v_ref.backward()
    x = a.clone().requires_grad_(True)
    v = fs.fused_ssim(x, b, padding)
    v.backward()
    assert abs(float(v) - float(v_ref)) < 1e-6
    assert rel_err(x.grad.cpu().numpy(), x_ref.grad.cpu().numpy()) < 1e-4
    # maps and derivative maps element-wise
    m_ref, d1, d2, d3 = ref.fusedssim(C1, C2, a, b, True)
    m = fs.FusedSSIMMap.apply(C1, C2, a, b, "same", False)
    assert float((m - m_ref).abs().max()) < MAP_TOL
    from gaussian_splatting_b200 import ssim as impl
    _, _, dms, _, _ = impl._forward(C1, C2, a, b, 0, True, False, False)
    for got, want in zip(dms, (d1, d2, d3)):
        assert rel_err(got.cpu().numpy(), want.cpu().numpy()) < 1e-4


def test_properties_at_full_size():
    import fused_ssim as fs
    a, b = _pair((1, 3, 1080, 1920), seed=11)
    # identical images: SSIM == 1 and the gradient vanishes
    x = a.clone().requires_grad_(True)
    v = fs.fused_ssim(x, a)
    v.backward()
    assert abs(float(v) - 1.0) < 1e-6 and float(x.grad.abs().max()) < 1e-9
    # deterministic: two runs are bit-identical (fixed-order reduction, no atomics)
    outs = []
    for _ in range(2):
        x = a.clone().requires_grad_(True)
        v = fs.fused_ssim(x, b)
        v.backward()
        outs.append((v.detach().clone(), x.grad.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    # the map is symmetric in its arguments; batch entries are independent
    m_ab = fs.FusedSSIMMap.apply(C1, C2, a, b, "same", False)
    m_ba = fs.FusedSSIMMap.apply(C1, C2, b, a, "same", False)
    assert float((m_ab - m_ba).abs().max()) < 1e-6
    two = fs.FusedSSIMMap.apply(C1, C2, torch.cat((a, b)), torch.cat((b, b)), "same", False)
    assert torch.equal(two[0], m_ab[0]) and float((two[1] - 1).abs().max()) < 1e-6
    assert float(m_ab.max()) <= 1 + 1e-6
    # linearity of the backward in the upstream gradient
    g = torch.Generator().manual_seed(2)
    d1, d2 = torch.randn(a.shape, generator=g).cuda(), torch.randn(a.shape, generator=g).cuda()

    def bwd(dl):
        x = a.clone().requires_grad_(True)
        fs.FusedSSIMMap.apply(C1, C2, x, b, "same", True).backward(dl)
        return x.grad
    assert rel_err((bwd(d1) + 2 * bwd(d2)).cpu().numpy(), bwd(d1 + 2 * d2).cpu().numpy()) < 1e-5


def test_host_behaviour():
    import fused_ssim as fs
    a, b = _pair((1, 3, 40, 50), seed=3)
    # non-contiguous inputs are accepted (the reference calls .contiguous())
    at = a.permute(0, 1, 3, 2)
    v = fs.fused_ssim(at, b.permute(0, 1, 3, 2))
    v_ref, _ = ssim_ref.fused_ssim(at.cpu().numpy(), b.permute(0, 1, 3, 2).cpu().numpy(), dtype=np.float64)
    assert abs(float(v) - float(v_ref)) < MAP_TOL
    # inference mode saves nothing; differentiating through it fails loudly
    x = a.clone().requires_grad_(True)
    v = fs.fused_ssim(x, b, train=False)
    with pytest.raises(RuntimeError, match="train=True"):
        v.backward()
    with pytest.raises(RuntimeError, match="equal shape"):
        fs.fused_ssim(a, b[:, :, :-1])
    with pytest.raises(RuntimeError, match="float32"):
        fs.fused_ssim(a.double(), b.double())
    # empty batch and empty crop
    e = torch.empty(0, 3, 16, 16, device="cuda")
    assert torch.isnan(fs.fused_ssim(e, e))
    s = torch.rand(1, 1, 8, 8, device="cuda")
    assert torch.isnan(fs.fused_ssim(s, s, "valid"))
    # runs on the caller's stream
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        v2 = fs.fused_ssim(a, b)
    st.synchronize()
    assert torch.equal(v2, fs.fused_ssim(a, b))


def test_ssim_kernels_are_what_ran():
    """gpu_launches evidence: the library's launch counter moves by 2 per fused_ssim forward
    (stencil + mean finish) and 1 per backward."""
    import fused_ssim as fs
    from gaussian_splatting_b200 import _lib
    a, b = _pair((1, 3, 64, 64), seed=5)
    x = a.clone().requires_grad_(True)
    n0 = _lib.launch_count()
    v = fs.fused_ssim(x, b)
    n1 = _lib.launch_count()
    v.backward()
    n2 = _lib.launch_count()
    assert (n1 - n0, n2 - n1) == (2, 1)
