"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/gsrast.h declares, struct layouts agree between the header and the ctypes mirror, the
Python surface matches the reference's (names, field order, messages), and nothing in the
product imports the oracle.  No compute entry point is called (there is no GPU here)."""
import ctypes as C
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from gaussian_splatting_b200 import _lib, build
    build.build()
    return _lib.load()


def _header():
    return open(os.path.join(ROOT, "include", "gsrast.h")).read()


def test_exports_every_declared_symbol(lib):
    from gaussian_splatting_b200 import _lib
    declared = set(re.findall(r"\b(gsr_[a-z0-9_]+)\s*\(", _header()))
    assert declared == set(_lib.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name


def test_version_and_error_string(lib):
    assert lib.gsr_version() == int(re.search(r"#define GSR_VERSION (\d+)", _header()).group(1))
    assert isinstance(lib.gsr_last_error(), bytes)


def test_struct_layouts_match_header():
    from gaussian_splatting_b200 import _lib
    h = _header()
    for name, cls in (("GsrParams", _lib.GsrParams), ("GsrBuffers", _lib.GsrBuffers),
                      ("GsrForwardOut", _lib.GsrForwardOut), ("GsrBackwardIO", _lib.GsrBackwardIO),
                      ("GsrSsimParams", _lib.GsrSsimParams), ("GsrLossParams", _lib.GsrLossParams),
                      ("GsrDensifyPlanArgs", _lib.GsrDensifyPlanArgs),
                      ("GsrDensifyTensor", _lib.GsrDensifyTensor), ("GsrDensifyApplyArgs", _lib.GsrDensifyApplyArgs)):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (name, name), h, re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        fields = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            names = decl.replace("*", " ").split(",")
            first = names[0].split()[-1]
            fields.append(first)
            fields += [n.strip() for n in names[1:]]
        fields = [f.split("[")[0] for f in fields]          # array members: name[COUNT]
        assert fields == [f[0] for f in cls._fields_], name
    assert C.sizeof(_lib.GsrParams) == 12 * 4 + 12 * 8
    assert C.sizeof(_lib.GsrBuffers) == 11 * 8          # + accumulator flags
    assert C.sizeof(_lib.GsrSsimParams) == 2 * 8 + 8 * 4        # 7 ints/floats + tail padding
    assert C.sizeof(_lib.GsrLossParams) == 5 * 8 + 8 * 4
    assert C.sizeof(_lib.GsrDensifyPlanArgs) == 8 * 4 + 5 * 8
    assert C.sizeof(_lib.GsrDensifyTensor) == 24
    n = int(re.search(r"#define GSR_DENSIFY_MAX_TENSORS (\d+)", h).group(1))
    assert n == _lib.DENSIFY_MAX_TENSORS and C.sizeof(_lib.GsrDensifyApplyArgs) == 6 * 4 + 3 * 8 + 8 + 24 * n
    for k in ("COPY", "MOMENT", "XYZ", "SCALING"):
        assert int(re.search(r"#define GSR_DENSIFY_%s\s+(\d+)" % k, h).group(1)) == getattr(_lib, "DENSIFY_" + k)


def test_workspace_size_queries(lib):
    # pure host arithmetic: monotone, 256-byte granular, >= the algorithmic minimum
    g1, g2 = lib.gsr_geom_bytes(1000), lib.gsr_geom_bytes(2000)
    assert g1 % 256 == 0 and g2 > g1 >= 1000 * 64
    i = lib.gsr_img_bytes(1920, 1080)
    assert i >= 1920 * 1080 * 8 + 8160 * 12
    assert lib.gsr_bin_bytes(0) > 0 and lib.gsr_bin_bytes(10 ** 6) >= 4 * 10 ** 6
    # the scratch arena has no num_rendered-sized part (the instances are never materialised outside `bin`);
    # it grows with the Gaussian count and with the tile grid
    s0 = lib.gsr_scratch_bytes(1000, 64, 64, 0)
    assert lib.gsr_scratch_bytes(1000, 64, 64, 50000) == s0 > 1000 * (8 + 8 + 8 + 8 + 48)
    assert lib.gsr_scratch_bytes(100000, 64, 64, 0) > s0 and lib.gsr_scratch_bytes(1000, 1920, 1080, 0) > s0
    # C3: 3M Gaussians at 1080p stay below 0.5 GB (it was 0.73 GB with the instance sort)
    assert lib.gsr_scratch_bytes(3_000_000, 1920, 1080, 0) < 500e6
    assert lib.gsr_flat_grad_floats(16) == 59               # 59 floats / Gaussian


def test_python_surface_matches_reference():
    import diff_gaussian_rasterization as dgr
    # DGR/diff_gaussian_rasterization/__init__.py:151-180 field order
    assert dgr.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix",
        "projmatrix", "sh_degree", "campos", "prefiltered", "debug", "antialiasing")
    for name in ("GaussianRasterizer", "rasterize_gaussians", "_RasterizeGaussians", "cpu_deep_copy_tuple"):
        assert hasattr(dgr, name)
    # train.py:41-45 / gaussian_model.py:246-251 probe for it and build it as (params, lr=, eps=); step(visible, N)
    from diff_gaussian_rasterization import SparseGaussianAdam
    import inspect
    assert issubclass(SparseGaussianAdam, torch.optim.Adam)
    assert list(inspect.signature(SparseGaussianAdam.__init__).parameters) == ["self", "params", "lr", "eps"]
    assert list(inspect.signature(SparseGaussianAdam.step).parameters)[:3] == ["self", "visibility", "N"]
    # the fork's positional order of GaussianRasterizer.forward (__init__.py:218) with `dc` and `raw_inputs` appended
    assert list(inspect.signature(dgr.GaussianRasterizer.forward).parameters) == [
        "self", "means3D", "means2D", "opacities", "shs", "colors_precomp", "scales", "rotations", "cov3D_precomp", "dc",
        "raw_inputs"]


def test_fused_ssim_surface_matches_reference():
    """submodules/fused-ssim/fused_ssim/__init__.py: names, signatures, padding modes, no CPU path."""
    import inspect
    import fused_ssim as fs
    assert fs.allowed_padding == ["same", "valid"]
    assert list(inspect.signature(fs.fused_ssim).parameters) == ["img1", "img2", "padding", "train"]
    assert inspect.signature(fs.fused_ssim).parameters["padding"].default == "same"
    assert inspect.signature(fs.fused_ssim).parameters["train"].default is True
    assert list(inspect.signature(fs.FusedSSIMMap.forward).parameters) == [
        "ctx", "C1", "C2", "img1", "img2", "padding", "train"]
    with pytest.raises(AssertionError):
        fs.fused_ssim(torch.zeros(1, 3, 16, 16), torch.zeros(1, 3, 16, 16), padding="reflect")
    with pytest.raises(RuntimeError, match="CUDA"):
        fs.fused_ssim(torch.zeros(1, 3, 16, 16), torch.zeros(1, 3, 16, 16))


def test_ssim_scratch_query(lib):
    assert lib.gsr_ssim_partials(1, 3, 1080, 1920) == 3 * 34 * 60
    assert lib.gsr_ssim_partials(2, 3, 1, 1) == 6 and lib.gsr_ssim_partials(0, 3, 8, 8) == 0


def _settings():
    import diff_gaussian_rasterization as dgr
    return dgr.GaussianRasterizationSettings(
        image_height=32, image_width=32, tanfovx=0.5, tanfovy=0.5, bg=torch.zeros(3), scale_modifier=1.0,
        viewmatrix=torch.eye(4), projmatrix=torch.eye(4), sh_degree=3, campos=torch.zeros(3),
        prefiltered=False, debug=False, antialiasing=False)


def test_argument_validation_messages():
    import diff_gaussian_rasterization as dgr
    r = dgr.GaussianRasterizer(_settings())
    m3, m2, op = torch.zeros(4, 3), torch.zeros(4, 3), torch.zeros(4, 1)
    sh, col = torch.zeros(4, 16, 3), torch.zeros(4, 3)
    sc, rot, cov = torch.ones(4, 3), torch.zeros(4, 4), torch.zeros(4, 6)
    with pytest.raises(Exception, match="Please provide exactly one of either SHs or precomputed colors!"):
        r(m3, m2, op, shs=None, colors_precomp=None, scales=sc, rotations=rot)
    with pytest.raises(Exception, match="Please provide exactly one of either SHs or precomputed colors!"):
        r(m3, m2, op, shs=sh, colors_precomp=col, scales=sc, rotations=rot)
    msg = "Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!"
    with pytest.raises(Exception, match=msg):
        r(m3, m2, op, shs=sh, scales=sc, rotations=None)
    with pytest.raises(Exception, match=msg):
        r(m3, m2, op, shs=sh, scales=sc, rotations=rot, cov3D_precomp=cov)


def test_no_cpu_fallback():
    """CPU tensors must fail loudly — the product has no CPU path."""
    import diff_gaussian_rasterization as dgr
    r = dgr.GaussianRasterizer(_settings())
    with pytest.raises(RuntimeError, match="CUDA"):
        r(torch.zeros(4, 3), torch.zeros(4, 3), torch.zeros(4, 1), shs=torch.zeros(4, 16, 3),
          scales=torch.ones(4, 3), rotations=torch.zeros(4, 4))
    with pytest.raises(RuntimeError, match="CUDA"):
        r.markVisible(torch.zeros(4, 3))
    with pytest.raises(RuntimeError, match=r"means3D must have dimensions \(num_points, 3\)"):
        r(torch.zeros(4, 2), torch.zeros(4, 3), torch.zeros(4, 1), shs=torch.zeros(4, 16, 3),
          scales=torch.ones(4, 3), rotations=torch.zeros(4, 4))


def test_product_never_touches_the_oracle():
    pkg_dirs = [os.path.join(ROOT, "gaussian_splatting_b200"), os.path.join(ROOT, "diff_gaussian_rasterization"),
                os.path.join(ROOT, "fused_ssim")]
    for d in pkg_dirs:
        for base, _, files in os.walk(d):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h")):
                    src = open(os.path.join(base, f)).read()
                    assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f
                    assert "libgsoracle" not in src and "gs_oracle" not in src, f


def test_cpu_deep_copy_tuple():
    import diff_gaussian_rasterization as dgr
    t = torch.arange(3.0)
    out = dgr.cpu_deep_copy_tuple((t, 5, "x"))
    assert out[1] == 5 and out[2] == "x" and out[0] is not t and torch.equal(out[0], t)
