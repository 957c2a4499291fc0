"""GPU check of density control against the reference's FORMULATION evaluated by torch on the same device
(profiles/bench_densify.py::torch_formulation: clone -> split -> prune with boolean masks, torch.cat and index, the op
sequence of scene/gaussian_model.py:400-621).  The numpy oracle runs on the host libm; this one shares the device's
exp / log / sigmoid with the kernels, so decisions must agree row for row at 200K Gaussians.  Sorted after every
other GPU test."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles"))


def test_densify_and_prune_equals_the_torch_formulation_on_the_device():
    import bench_densify as B
    from gaussian_splatting_b200 import densify
    dev = torch.device("cuda", 0)
    s = B.make_state(200_000, dev, seed=4)
    gm = B.model_from(s)
    info = densify.densify_and_prune(gm, 0.0002, 0.005, 5.0, 20, None, noise=s["noise"])
    want = B.torch_formulation(s)
    assert info["clones"] > 1000 and info["split"] > 1000
    for n in B.GROUPS:
        p = getattr(gm, B.ATTR[n]).detach()
        st = gm.optimizer.state[getattr(gm, B.ATTR[n])]
        assert p.shape == want[n].shape, (n, tuple(p.shape), tuple(want[n].shape))
        if n in ("xyz", "scaling"):
            assert torch.allclose(p, want[n], rtol=1e-5, atol=1e-5), n
        else:
            assert torch.equal(p, want[n]), n
        assert torch.equal(st["exp_avg"], want["m_" + n]) and torch.equal(st["exp_avg_sq"], want["v_" + n]), n
