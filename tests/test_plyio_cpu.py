"""CPU tests of the Gaussian file formats: gaussian_splatting_b200/plyio.py against
oracle/ply_ref.py (the reference's save_ply / load_ply restated with struct and loops, scene/gaussian_model.py:285-380).
The product's two kernels (gsr_ply_pack / gsr_ply_unpack) are served here by the CPU emulation of their per-element
function (tests/host_emul, monkeypatched in this test only); tests/test_zz_plyio_gpu.py runs the real ones."""
import contextlib
import types

import numpy as np
import pytest
import torch
from torch import nn

from gaussian_splatting_b200 import _lib, plyio
from oracle import ply_ref
from tests.test_densify_emul_cpu import FakeLib, emul, kernel_emul, kernel_emul_wide  # noqa: F401  (fixtures)


@pytest.fixture(params=("per-element", "kernels", "kernels-64bit-index"))
def on_cpu(request, emul, kernel_emul, kernel_emul_wide, monkeypatch):  # noqa: F811
    fake = {"per-element": FakeLib(emul), "kernels": kernel_emul, "kernels-64bit-index": kernel_emul_wide}[request.param]
    monkeypatch.setattr(_lib, "load", lambda: fake)
    monkeypatch.setattr(plyio, "_require_cuda", lambda t, what: None)
    monkeypatch.setattr(plyio, "_device_ctx", lambda d: contextlib.nullcontext())
    monkeypatch.setattr(plyio, "_stream", lambda d: None)
    monkeypatch.setattr(plyio, "_to_host", lambda rows: rows)
    monkeypatch.setattr(plyio, "_to_device", lambda host, device: host)
    return fake


def model(P, degree, seed=0):
    g = torch.Generator().manual_seed(seed)
    K = (degree + 1) ** 2 - 1
    gm = types.SimpleNamespace(max_sh_degree=degree, active_sh_degree=0)
    gm._xyz = nn.Parameter(torch.randn(P, 3, generator=g))
    gm._features_dc = nn.Parameter(torch.randn(P, 1, 3, generator=g))
    gm._features_rest = nn.Parameter(torch.randn(P, K, 3, generator=g))
    gm._opacity = nn.Parameter(torch.randn(P, 1, generator=g))
    gm._scaling = nn.Parameter(torch.randn(P, 3, generator=g))
    gm._rotation = nn.Parameter(torch.randn(P, 4, generator=g))
    return gm


def arrays(gm):
    return [getattr(gm, a).detach().numpy() for a in ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling",
                                                      "_rotation")]


@pytest.mark.parametrize("P,degree", ((1, 3), (37, 3), (64, 0), (19, 1), (0, 3)))
def test_save_ply_bytes_match_the_oracle(on_cpu, tmp_path, P, degree):
    gm = model(P, degree, seed=P)
    path = str(tmp_path / "point_cloud" / "iteration_7" / "point_cloud.ply")       # scene/__init__.py:101-102 layout
    plyio.save_ply(gm, path)
    assert open(path, "rb").read() == ply_ref.save_ply_bytes(*arrays(gm))
    assert plyio.construct_list_of_attributes(gm) == ply_ref.attribute_names(3, 3 * ((degree + 1) ** 2 - 1))


@pytest.mark.parametrize("P,degree", ((1, 3), (37, 3), (64, 0), (19, 2)))
def test_load_ply_matches_the_oracle_and_round_trips(on_cpu, tmp_path, P, degree):
    gm = model(P, degree, seed=100 + P)
    data = ply_ref.save_ply_bytes(*arrays(gm))
    path = str(tmp_path / "in.ply")
    open(path, "wb").write(data)
    got = types.SimpleNamespace(max_sh_degree=degree, active_sh_degree=0)
    plyio.load_ply(got, path, device="cpu")
    want = ply_ref.load_ply_bytes(data, degree)
    for attr, key in (("_xyz", "xyz"), ("_features_dc", "f_dc"), ("_features_rest", "f_rest"), ("_opacity", "opacity"),
                      ("_scaling", "scaling"), ("_rotation", "rotation")):
        p = getattr(got, attr)
        assert isinstance(p, nn.Parameter) and p.requires_grad and p.is_contiguous()
        assert np.array_equal(p.detach().numpy(), want[key]), attr
        assert np.array_equal(p.detach().numpy(), getattr(gm, attr).detach().numpy()), attr      # round trip
    assert got.active_sh_degree == degree
    with pytest.raises(AssertionError):                       # gaussian_model.py:353
        plyio.load_ply(types.SimpleNamespace(max_sh_degree=degree + 1), path, device="cpu")


def test_load_ply_reads_by_name_any_order_type_and_encoding(on_cpu, tmp_path):
    gm = model(11, 1, seed=3)
    names = plyio.construct_list_of_attributes(gm)
    rows = plyio.pack_rows(gm).numpy()
    want = [a.copy() for a in arrays(gm)]
    rng = np.random.default_rng(0)
    perm = rng.permutation(len(names))
    # (a) shuffled property order, doubles, an extra property, big endian
    dt = np.dtype([(names[i], ">f8") for i in perm] + [("extra", ">i4")])
    el = np.zeros(11, dt)
    for i in perm:
        el[names[i]] = rows[:, i]
    head = "ply\nformat binary_big_endian 1.0\ncomment made by a test\nelement vertex 11\n"
    head += "".join(f"property float64 {names[i]}\n" for i in perm) + "property int extra\n"
    head += "element face 0\nproperty list uchar int vertex_indices\nend_header\n"
    pa = str(tmp_path / "a.ply")
    open(pa, "wb").write(head.encode() + el.tobytes())
    # (b) ascii
    head = "ply\nformat ascii 1.0\nelement vertex 11\n" + "".join(f"property float {n}\n" for n in names) + "end_header\n"
    pb = str(tmp_path / "b.ply")
    open(pb, "w").write(head + "\n".join(" ".join(repr(float(v)) for v in r) for r in rows) + "\n")
    for path in (pa, pb):
        got = types.SimpleNamespace(max_sh_degree=1)
        plyio.load_ply(got, path, device="cpu")
        for a, w in zip(arrays(got), want):
            assert np.array_equal(a, w)
    for bad in (b"plx\n", b"ply\nformat binary_little_endian 1.0\nelement vertex 1\nproperty list uchar int x\nend_header\n",
                b"ply\nformat binary_little_endian 1.0\nelement vertex 2\nproperty float x\nend_header\n\0\0\0\0"):
        pc = str(tmp_path / "c.ply")
        open(pc, "wb").write(bad)
        with pytest.raises(ValueError):
            plyio.read_ply_table(pc)


def test_point_cloud_ply_round_trip(tmp_path):
    # storePly / fetchPly (scene/dataset_readers.py:120-143): host-only, no kernels
    rng = np.random.default_rng(1)
    xyz = rng.normal(size=(50, 3)).astype(np.float32)
    rgb = rng.integers(0, 256, size=(50, 3)).astype(np.uint8)
    path = str(tmp_path / "points3D.ply")
    plyio.storePly(path, xyz, rgb)
    raw = open(path, "rb").read()
    head = ("ply\nformat binary_little_endian 1.0\nelement vertex 50\nproperty float x\nproperty float y\nproperty float z\n"
            "property float nx\nproperty float ny\nproperty float nz\nproperty uchar red\nproperty uchar green\n"
            "property uchar blue\nend_header\n").encode()
    assert raw.startswith(head) and len(raw) == len(head) + 50 * 27
    pos, col, nrm = plyio.fetchPly(path)
    assert np.array_equal(pos, xyz) and np.array_equal(col, rgb / 255.0) and not nrm.any()


def test_checkpoint_round_trip(tmp_path):
    # train.py:256 / :62 — (capture(), iteration) through torch.save / torch.load
    gm = model(5, 1)
    gm.max_radii2D, gm.xyz_gradient_accum, gm.denom = torch.zeros(5), torch.ones(5, 1), torch.ones(5, 1) * 2
    gm.spatial_lr_scale = 1.5
    gm.optimizer = torch.optim.Adam([{"params": [gm._xyz], "lr": 1e-3, "name": "xyz"}], lr=0.0, eps=1e-15)
    gm._xyz.grad = torch.ones_like(gm._xyz)
    gm.optimizer.step()
    path = str(tmp_path / "chkpnt7000.pth")
    plyio.save_checkpoint(gm, 7000, path)
    params, it = plyio.load_checkpoint(path)
    assert it == 7000 and len(params) == 12 and params[0] == 0 and params[11] == 1.5
    assert torch.equal(params[1], gm._xyz) and torch.equal(params[9], gm.denom)
    assert params[10]["param_groups"][0]["name"] == "xyz" and float(params[10]["state"][0]["step"]) == 1.0


@pytest.mark.skipif(not __import__("os").path.isdir("/root/reference/scene"), reason="needs the reference tree")
def test_reference_save_and_load_ply_through_plyfile_compat(on_cpu, tmp_path):  # noqa: F811
    """The REFERENCE's own save_ply / load_ply / storePly / fetchPly (unmodified) running on plyfile_compat: its row
    assembly (gaussian_model.py:303-321) must give the bytes plyio.save_ply and the oracle give, and its load_ply
    (:329-380) must read back what plyio.load_ply reads."""
    import importlib.util
    import os
    import sys
    from gaussian_splatting_b200 import plyfile_compat
    saved = {n: getattr(torch, n) for n in ("zeros", "ones", "eye", "tensor", "empty", "normal")}
    names = ("plyfile", "simple_knn", "simple_knn._C", "scene", "scene.gaussian_model", "scene.dataset_readers", "utils",
             "utils.general_utils", "utils.system_utils", "utils.sh_utils", "utils.graphics_utils")
    saved_mods = {k: sys.modules.get(k) for k in names}
    saved_path = list(sys.path)
    try:
        sys.modules.pop("plyfile", None)
        assert plyfile_compat.install() is plyfile_compat
        here = os.path.dirname(os.path.abspath(__file__))
        spec = importlib.util.spec_from_file_location("mk_golden", os.path.join(here, "golden", "make_densify_golden.py"))
        mk = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mk)                       # device="cuda" redirect + the reference's GaussianModel
        gm = model(23, 3, seed=5)
        ref = mk.GaussianModel(3)
        for a in ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation"):
            setattr(ref, a, getattr(gm, a))
        p_ref, p_ours = str(tmp_path / "ref" / "pc.ply"), str(tmp_path / "ours" / "pc.ply")
        ref.save_ply(p_ref)                               # the reference's code, compat writer
        plyio.save_ply(gm, p_ours)
        data = open(p_ref, "rb").read()
        assert data == open(p_ours, "rb").read() == ply_ref.save_ply_bytes(*arrays(gm))
        back = mk.GaussianModel(3)
        back.load_ply(p_ours)                             # the reference's code, compat reader
        ours = types.SimpleNamespace(max_sh_degree=3)
        plyio.load_ply(ours, p_ref, device="cpu")
        for a in ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation"):
            assert torch.equal(getattr(back, a).detach(), getattr(ours, a).detach()), a
            assert torch.equal(getattr(back, a).detach(), getattr(gm, a).detach()), a
        assert back.active_sh_degree == ours.active_sh_degree == 3
    finally:
        for n, f in saved.items():
            setattr(torch, n, f)
        sys.path[:] = saved_path
        for k, m in saved_mods.items():
            if m is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = m


def test_plyfile_compat_api(tmp_path):
    from gaussian_splatting_b200.plyfile_compat import PlyData, PlyElement
    dtype = [("x", "f4"), ("y", "f4"), ("z", "f4"), ("red", "u1")]
    el = np.zeros(3, dtype)
    el["x"], el["red"] = [1, 2, 3], [7, 8, 9]
    path = str(tmp_path / "t.ply")
    PlyData([PlyElement.describe(el, "vertex")]).write(path)
    back = PlyData.read(path)
    assert [p.name for p in back.elements[0].properties] == ["x", "y", "z", "red"]
    assert np.array_equal(back["vertex"]["x"], el["x"]) and np.array_equal(back.elements[0]["red"], el["red"])
    assert len(back.elements[0]) == 3
    with pytest.raises(KeyError):
        back["face"]
    with pytest.raises(TypeError):
        PlyElement.describe(np.zeros((3, 2)), "vertex")
