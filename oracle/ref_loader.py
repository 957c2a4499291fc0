"""Imports the reference's own CUDA rasterizer built by oracle/build_ref.py (oracle/_ref/).

TEST / BASELINE INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), bench.py --impl reference).
The reference package is called `diff_gaussian_rasterization`, the same name as the drop-in, so
it is imported under a private module name (`_gsref_<variant>`); its `from . import _C` then
resolves to oracle/_ref/<variant>/diff_gaussian_rasterization/_C.so.
"""
import importlib.util
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
_cache = {}


def available(variant="restored"):
    return os.path.exists(os.path.join(_HERE, "_ref", variant, "diff_gaussian_rasterization", "_C.so"))


def load(variant="restored"):
    """Returns the reference package (GaussianRasterizationSettings, GaussianRasterizer, _C)."""
    if variant in _cache:
        return _cache[variant]
    pkg_dir = os.path.join(_HERE, "_ref", variant, "diff_gaussian_rasterization")
    if not os.path.exists(os.path.join(pkg_dir, "_C.so")):
        raise RuntimeError(f"reference build '{variant}' missing: run `python oracle/build_ref.py` "
                           "in the authoring container (needs /root/reference)")
    import torch  # noqa: F401  (libtorch must be loaded before _C.so)
    name = f"_gsref_{variant}"
    spec = importlib.util.spec_from_file_location(name, os.path.join(pkg_dir, "__init__.py"),
                                                  submodule_search_locations=[pkg_dir])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    _cache[variant] = mod
    return mod


def fused_ssim_available():
    return os.path.exists(os.path.join(_HERE, "_ref", "fused_ssim", "fused_ssim_cuda.so"))


def load_fused_ssim():
    """The reference's fused_ssim package (fused_ssim(), FusedSSIMMap) over its own ssim.cu built for sm_100.
    Its wrapper does `from fused_ssim_cuda import ...`, so the extension is registered under that name; the
    package itself is imported under a private name (the drop-in is also called `fused_ssim`)."""
    if "fused_ssim" in _cache:
        return _cache["fused_ssim"]
    d = os.path.join(_HERE, "_ref", "fused_ssim")
    so = os.path.join(d, "fused_ssim_cuda.so")
    if not os.path.exists(so):
        raise RuntimeError("reference fused-ssim build missing: run `python oracle/build_ref.py ssim` "
                           "in the authoring container (needs /root/reference)")
    import torch  # noqa: F401
    spec = importlib.util.spec_from_file_location("fused_ssim_cuda", so)
    ext = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ext)
    sys.modules["fused_ssim_cuda"] = ext
    spec = importlib.util.spec_from_file_location("_gsref_fused_ssim", os.path.join(d, "fused_ssim", "__init__.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["_gsref_fused_ssim"] = mod
    spec.loader.exec_module(mod)
    _cache["fused_ssim"] = mod
    return mod


def forward_state(ref, settings, means3D, sh, colors_precomp, opacities, scales, rotations, cov3D_precomp):
    """Calls the reference's native forward directly and decodes its state buffers
    (layout: rasterizer_impl.cu:184-223 fromChunk order, 128-byte aligned `obtain`).

    Returns dict(num_rendered, color, radii, invdepth, point_list, point_list_keys, ranges,
    n_contrib, final_T, geomBuffer, binningBuffer, imgBuffer)."""
    import torch
    e = torch.Tensor([])
    args = (settings.bg, means3D, colors_precomp if colors_precomp is not None else e, opacities,
            scales if scales is not None else e, rotations if rotations is not None else e,
            settings.scale_modifier, cov3D_precomp if cov3D_precomp is not None else e,
            settings.viewmatrix, settings.projmatrix, settings.tanfovx, settings.tanfovy,
            settings.image_height, settings.image_width, sh if sh is not None else e, settings.sh_degree,
            settings.campos, settings.prefiltered, settings.antialiasing, settings.debug)
    R, color, radii, geom, binning, img, invdepth = ref._C.rasterize_gaussians(*args)
    H, W = settings.image_height, settings.image_width
    N = H * W

    def carve(buf, specs):
        out, off = {}, 0
        base = buf.data_ptr()
        for nm, dtype, count in specs:
            itemsize = torch.empty(0, dtype=dtype).element_size()
            addr = (base + off + 127) // 128 * 128
            off = addr - base
            out[nm] = buf[off:off + count * itemsize].view(dtype)
            off += count * itemsize
        return out

    b = carve(binning, [("point_list", torch.int32, R), ("point_list_unsorted", torch.int32, R),
                        ("point_list_keys", torch.int64, R), ("point_list_keys_unsorted", torch.int64, R)])
    im = carve(img, [("accum_alpha", torch.float32, N), ("n_contrib", torch.int32, N),
                     ("ranges", torch.int32, 2 * N)])
    T = ((W + 15) // 16) * ((H + 15) // 16)
    return dict(num_rendered=R, color=color, radii=radii, invdepth=invdepth,
                point_list=b["point_list"], point_list_keys=b["point_list_keys"],
                ranges=im["ranges"][:2 * T].view(T, 2), n_contrib=im["n_contrib"], final_T=im["accum_alpha"],
                geomBuffer=geom, binningBuffer=binning, imgBuffer=img)
