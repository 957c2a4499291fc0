"""CPU oracle for the differentiable Gaussian rasterizer (numpy + ctypes over libgsoracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never by gaussian_splatting_b200/.

The C sources restate the reference's CUDA kernels function by function (citations in
gs_oracle*.c).  This module only marshals numpy arrays; the stage order is that of
CudaRasterizer::Rasterizer::forward / ::backward
(submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer_impl.cu:227-370, 374-479).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

_f = np.float32


def build(force=False):
    """Compile libgsoracle.so with gcc (oracle/Makefile) if missing or stale."""
    so = os.path.join(_HERE, "libgsoracle.so")
    srcs = [os.path.join(_HERE, n) for n in
            ("gs_oracle.c", "gs_oracle_prebwd.c", "gs_oracle_render.c", "gso_parallel.h", "Makefile")]
    stale = force or not os.path.exists(so) or any(
        os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if stale and all(os.path.exists(s) for s in srcs):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "libgsoracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.gso_binning.restype = C.c_int64
        _LIB.gso_higher_msb.restype = C.c_uint32
        _LIB.gso_get_threads.restype = C.c_int
    return _LIB


def set_threads(n):
    lib().gso_set_threads(C.c_int(int(n)))


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _arr(a, dtype=_f):
    if a is None:
        return None
    a = np.ascontiguousarray(np.asarray(a), dtype=dtype)
    return a if a.size else None


def higher_msb(n):
    return int(lib().gso_higher_msb(C.c_uint32(n)))


def mark_visible(means3D, viewmatrix):
    means3D = _arr(means3D)
    P = means3D.shape[0]
    out = np.zeros(P, np.uint8)
    lib().gso_mark_visible(C.c_int(P), _p(means3D), _p(_arr(viewmatrix)), _p(out))
    return out.astype(bool)


def preprocess(inp, s):
    """inp: dict(means3D, opacities, shs|colors_precomp, scales+rotations|cov3D_precomp);
    s: settings dict(image_height, image_width, tanfovx, tanfovy, bg, scale_modifier,
    viewmatrix, projmatrix, sh_degree, campos, antialiasing)."""
    means3D = _arr(inp["means3D"])
    P = means3D.shape[0]
    shs = _arr(inp.get("shs"))
    M = shs.shape[1] if shs is not None else 0
    st = dict(
        P=P, M=M,
        radii=np.zeros(P, np.int32), means2D=np.zeros((P, 2), _f), depths=np.zeros(P, _f),
        cov3D=np.zeros((P, 6), _f), rgb=np.zeros((P, 3), _f), clamped=np.zeros((P, 3), np.uint8),
        conic_opacity=np.zeros((P, 4), _f), tiles_touched=np.zeros(P, np.uint32))
    a = dict(means3D=means3D, scales=_arr(inp.get("scales")), rotations=_arr(inp.get("rotations")),
             opacities=_arr(inp["opacities"]), shs=shs, cov3D_precomp=_arr(inp.get("cov3D_precomp")),
             colors_precomp=_arr(inp.get("colors_precomp")), viewmatrix=_arr(s["viewmatrix"]),
             projmatrix=_arr(s["projmatrix"]), campos=_arr(s["campos"]))
    st["_in"] = a
    lib().gso_preprocess(
        C.c_int(P), C.c_int(int(s["sh_degree"])), C.c_int(M), _p(a["means3D"]), _p(a["scales"]),
        C.c_float(s["scale_modifier"]), _p(a["rotations"]), _p(a["opacities"]), _p(a["shs"]),
        _p(a["cov3D_precomp"]), _p(a["colors_precomp"]), _p(a["viewmatrix"]), _p(a["projmatrix"]),
        _p(a["campos"]), C.c_int(int(s["image_width"])), C.c_int(int(s["image_height"])),
        C.c_float(s["tanfovx"]), C.c_float(s["tanfovy"]), C.c_int(int(bool(s["antialiasing"]))),
        _p(st["radii"]), _p(st["means2D"]), _p(st["depths"]), _p(st["cov3D"]), _p(st["rgb"]),
        _p(st["clamped"]), _p(st["conic_opacity"]), _p(st["tiles_touched"]))
    return st


def binning(W, H, radii, means2D, depths, tiles_touched):
    """Returns dict(R, keys (sorted u64), point_list (sorted u32), ranges (T,2))."""
    P = radii.shape[0]
    R = int(np.asarray(tiles_touched, np.uint64).sum())
    T = ((W + 15) // 16) * ((H + 15) // 16)
    n = max(R, 1)
    keys = np.zeros(n, np.uint64); vals = np.zeros(n, np.uint32)
    tk = np.zeros(n, np.uint64); tv = np.zeros(n, np.uint32)
    ranges = np.zeros((T, 2), np.uint32)
    R2 = lib().gso_binning(C.c_int(P), C.c_int(W), C.c_int(H), _p(_arr(radii, np.int32)),
                           _p(_arr(means2D)), _p(_arr(depths)), _p(_arr(tiles_touched, np.uint32)),
                           _p(keys), _p(vals), _p(tk), _p(tv), _p(ranges), C.c_int64(n))
    assert R2 == R
    return dict(R=R, keys=keys[:R], point_list=vals[:R], ranges=ranges)


def render_forward(W, H, ranges, point_list, means2D, features, conic_opacity, depths, bg,
                   tile_lo=0, tile_hi=-1):
    N = W * H
    out = dict(final_T=np.zeros(N, _f), n_contrib=np.zeros(N, np.uint32),
               color=np.zeros((3, H, W), _f), invdepth=np.zeros((1, H, W), _f),
               stats=np.zeros(2, np.int64))
    pl = _arr(point_list, np.uint32)
    if pl is None:
        pl = np.zeros(1, np.uint32)
    lib().gso_render_forward(
        C.c_int(W), C.c_int(H), _p(_arr(ranges, np.uint32)), _p(pl), _p(_arr(means2D)),
        _p(_arr(features)), _p(_arr(conic_opacity)), _p(_arr(depths)), _p(_arr(bg)),
        _p(out["final_T"]), _p(out["n_contrib"]), _p(out["color"]), _p(out["invdepth"]),
        C.c_int(tile_lo), C.c_int(tile_hi), _p(out["stats"]))
    return out


def render_backward(P, W, H, ranges, point_list, bg, means2D, conic_opacity, colors, depths, final_T,
                    n_contrib, dL_dpix, dL_dinvdepth_pix=None, tile_lo=0, tile_hi=-1):
    g = dict(dL_dmean2D=np.zeros((P, 3), _f), dL_dconic=np.zeros((P, 4), _f),
             dL_dopacity=np.zeros(P, _f), dL_dcolors=np.zeros((P, 3), _f),
             dL_dinvdepths=np.zeros(P, _f) if dL_dinvdepth_pix is not None else None)
    pl = _arr(point_list, np.uint32)
    if pl is None:
        pl = np.zeros(1, np.uint32)
    lib().gso_render_backward(
        C.c_int(P), C.c_int(W), C.c_int(H), _p(_arr(ranges, np.uint32)), _p(pl), _p(_arr(bg)),
        _p(_arr(means2D)), _p(_arr(conic_opacity)), _p(_arr(colors)), _p(_arr(depths)),
        _p(_arr(final_T)), _p(_arr(n_contrib, np.uint32)), _p(_arr(dL_dpix)), _p(_arr(dL_dinvdepth_pix)),
        _p(g["dL_dmean2D"]), _p(g["dL_dconic"]), _p(g["dL_dopacity"]), _p(g["dL_dcolors"]),
        _p(g["dL_dinvdepths"]), C.c_int(tile_lo), C.c_int(tile_hi))
    return g


def forward(inp, s, tile_lo=0, tile_hi=-1):
    """Full forward: returns dict with color (3,H,W), invdepth (1,H,W), radii, and all state."""
    W, H = int(s["image_width"]), int(s["image_height"])
    st = preprocess(inp, s)
    b = binning(W, H, st["radii"], st["means2D"], st["depths"], st["tiles_touched"])
    feats = st["_in"]["colors_precomp"] if st["_in"]["colors_precomp"] is not None else st["rgb"]
    r = render_forward(W, H, b["ranges"], b["point_list"], st["means2D"], feats, st["conic_opacity"],
                       st["depths"], s["bg"], tile_lo, tile_hi)
    out = dict(st)
    out.update(b)
    out.update(r)
    out["features"] = feats
    out["W"], out["H"] = W, H
    return out


def backward(fw, s, dL_dcolor, dL_dinvdepth=None, restored=True, tile_lo=0, tile_hi=-1):
    """Backward of forward(); restored=False keeps the fork's truncated computeCov2DCUDA (F2)."""
    P, M, W, H = fw["P"], fw["M"], fw["W"], fw["H"]
    a = fw["_in"]
    g = render_backward(P, W, H, fw["ranges"], fw["point_list"], s["bg"], fw["means2D"],
                        fw["conic_opacity"], fw["features"], fw["depths"], fw["final_T"], fw["n_contrib"],
                        dL_dcolor, dL_dinvdepth, tile_lo, tile_hi)
    out = dict(dL_dmeans2D=g["dL_dmean2D"], dL_dcolors=g["dL_dcolors"],
               dL_dopacity=g["dL_dopacity"].copy(), dL_dconic=g["dL_dconic"],
               dL_dmeans3D=np.zeros((P, 3), _f), dL_dcov3D=np.zeros((P, 6), _f),
               dL_dsh=np.zeros((P, max(M, 0), 3), _f), dL_dscales=np.zeros((P, 3), _f),
               dL_drotations=np.zeros((P, 4), _f))
    cov3D = a["cov3D_precomp"] if a["cov3D_precomp"] is not None else fw["cov3D"]
    lib().gso_preprocess_backward(
        C.c_int(P), C.c_int(int(s["sh_degree"])), C.c_int(M), _p(a["means3D"]), _p(fw["radii"]),
        _p(a["shs"]), _p(fw["clamped"]), _p(a["opacities"]), _p(a["scales"]), _p(a["rotations"]),
        C.c_float(s["scale_modifier"]), _p(_arr(cov3D)), _p(a["viewmatrix"]), _p(a["projmatrix"]),
        C.c_int(W), C.c_int(H), C.c_float(s["tanfovx"]), C.c_float(s["tanfovy"]), _p(a["campos"]),
        _p(out["dL_dmeans2D"]), _p(out["dL_dconic"]), _p(g["dL_dinvdepths"]), _p(out["dL_dopacity"]),
        _p(out["dL_dmeans3D"]), _p(out["dL_dcolors"]), _p(out["dL_dcov3D"]), _p(out["dL_dsh"]),
        _p(out["dL_dscales"]), _p(out["dL_drotations"]), C.c_int(int(bool(s["antialiasing"]))),
        C.c_int(int(bool(restored))))
    return out
