"""GPU tests of the Gaussian file formats: the real gsr_ply_pack / gsr_ply_unpack kernels through
gaussian_splatting_b200.plyio against oracle/ply_ref.py (small sizes, file bytes exact) and, at 1M Gaussians, through
the save -> load round trip and the column layout of the row matrix."""
import types

import numpy as np
import pytest
import torch
from torch import nn

from oracle import ply_ref

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ATTRS = ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation")


def model(P, degree, seed=0):
    g = torch.Generator().manual_seed(seed)
    K = (degree + 1) ** 2 - 1
    gm = types.SimpleNamespace(max_sh_degree=degree, active_sh_degree=0)
    for attr, shape in zip(ATTRS, ((P, 3), (P, 1, 3), (P, K, 3), (P, 1), (P, 3), (P, 4))):
        setattr(gm, attr, nn.Parameter(torch.randn(*shape, generator=g).to(DEV)))
    return gm


def arrays(gm):
    return [getattr(gm, a).detach().cpu().numpy() for a in ATTRS]


@pytest.mark.parametrize("P,degree", ((1, 3), (37, 3), (64, 0), (19, 1), (0, 3), (1000, 2)))
def test_save_ply_bytes_and_load_ply_match_the_oracle(tmp_path, P, degree):
    from gaussian_splatting_b200 import plyio
    gm = model(P, degree, seed=P)
    path = str(tmp_path / "point_cloud" / "iteration_7" / "point_cloud.ply")
    plyio.save_ply(gm, path)
    data = open(path, "rb").read()
    assert data == ply_ref.save_ply_bytes(*arrays(gm))
    got = types.SimpleNamespace(max_sh_degree=degree)
    plyio.load_ply(got, path, device=DEV)
    want = ply_ref.load_ply_bytes(data, degree)
    for attr, key in zip(ATTRS, ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")):
        p = getattr(got, attr)
        assert isinstance(p, nn.Parameter) and p.requires_grad and p.is_cuda
        assert np.array_equal(p.detach().cpu().numpy(), want[key]), attr
    assert got.active_sh_degree == degree


def test_non_contiguous_parameters_and_shuffled_file(tmp_path):
    from gaussian_splatting_b200 import plyio
    gm = model(300, 3, seed=9)
    gm._features_rest = nn.Parameter(gm._features_rest.detach().transpose(1, 2).contiguous().transpose(1, 2))
    assert not gm._features_rest.is_contiguous()
    path = str(tmp_path / "a.ply")
    plyio.save_ply(gm, path)
    assert open(path, "rb").read() == ply_ref.save_ply_bytes(*arrays(gm))
    # a file with another property order and doubles is read by name
    names = plyio.construct_list_of_attributes(gm)
    rows = plyio.pack_rows(gm).cpu().numpy()
    perm = np.random.default_rng(0).permutation(len(names))
    el = np.zeros(300, np.dtype([(names[i], "<f8") for i in perm]))
    for i in perm:
        el[names[i]] = rows[:, i]
    head = "ply\nformat binary_little_endian 1.0\nelement vertex 300\n" + "".join(f"property double {names[i]}\n" for i in perm)
    pb = str(tmp_path / "b.ply")
    open(pb, "wb").write((head + "end_header\n").encode() + el.tobytes())
    got = types.SimpleNamespace(max_sh_degree=3)
    plyio.load_ply(got, pb, device=DEV)
    for attr in ATTRS:
        assert torch.equal(getattr(got, attr).detach(), getattr(gm, attr).detach()), attr


def test_one_million_gaussians_round_trip(tmp_path):
    from gaussian_splatting_b200 import plyio
    P, K = 1_000_000, 15
    gm = model(P, 3, seed=1)
    rows = plyio.pack_rows(gm)
    assert rows.shape == (P, 62)
    # column layout (gaussian_model.py:285-321): f_rest_{c*K + k} = _features_rest[p, k, c]; normals are zero
    assert torch.equal(rows[:, 0:3], gm._xyz.detach()) and not bool(rows[:, 3:6].any())
    assert torch.equal(rows[:, 6:9], gm._features_dc.detach()[:, 0, :])
    assert torch.equal(rows[:, 9:54], gm._features_rest.detach().transpose(1, 2).flatten(start_dim=1))
    assert torch.equal(rows[:, 54:55], gm._opacity.detach()) and torch.equal(rows[:, 55:58], gm._scaling.detach())
    assert torch.equal(rows[:, 58:62], gm._rotation.detach())
    path = str(tmp_path / "big.ply")
    plyio.save_ply(gm, path)
    got = types.SimpleNamespace(max_sh_degree=3)
    plyio.load_ply(got, path, device=DEV)
    for attr in ATTRS:
        assert torch.equal(getattr(got, attr).detach(), getattr(gm, attr).detach()), attr


def test_cpu_tensors_are_refused(tmp_path):
    from gaussian_splatting_b200 import plyio
    gm = model(4, 0)
    gm._xyz = nn.Parameter(gm._xyz.detach().cpu())
    with pytest.raises(RuntimeError):
        plyio.save_ply(gm, str(tmp_path / "x.ply"))
