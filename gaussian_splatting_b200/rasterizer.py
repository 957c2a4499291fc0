"""Host-side mirror of the reference's autograd boundary over the C-ABI of libgsrast.so.

Same names, argument order, defaults, exception messages, output tuple and gradient tuple as
submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py
(GaussianRasterizationSettings :151-180, GaussianRasterizer :183-274, _RasterizeGaussians
:48-148, rasterize_gaussians :24-46, cpu_deep_copy_tuple :18-21), so that
gaussian_renderer/__init__.py:render() and train.py run unmodified.

PyTorch is plumbing only (device memory, streams, autograd registration): every kernel on the
path lives in libgsrast.so.  There is no CPU or eager fallback: non-CUDA tensors or a missing
library raise.
"""
import ctypes as C
import os
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _lib
from ._lib import GsrBackwardIO, GsrBuffers, GsrForwardOut, GsrParams


def cpu_deep_copy_tuple(input_tuple):
    copied_tensors = [item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple]
    return tuple(copied_tensors)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    antialiasing: bool


# ---------------------------------------------------------------------------------------
# Workspace: the C layer never allocates.  Per device we keep one scratch arena (grown
# geometrically, reused by every call on that device's current stream) and a pinned
# int32[2] that receives {num_rendered, overflow}.  State that backward needs (geom/img/bin)
# is allocated per call and kept alive by the autograd ctx, as the reference does with
# geomBuffer/binningBuffer/imgBuffer (__init__.py:93).
# ---------------------------------------------------------------------------------------
class _DeviceWorkspace:
    def __init__(self, device):
        self.device = device
        self.scratch = torch.empty(0, dtype=torch.uint8, device=device)
        self.pinned = torch.zeros(2, dtype=torch.int32).pin_memory()
        # recorded by stage 1 right after the copy into `pinned`; the host waits on it, not on the stream
        self.counted = torch.cuda.Event()
        self.counted.record(torch.cuda.current_stream(device))     # materialises the cudaEvent_t handle
        self.last_R = 0
        self.regrows = 0
        # (P, W, H) of the forward that left the backward's gradient accumulator (a part of `scratch`) zeroed, while
        # nothing else has used the arena since (GsrBuffers.accumulator, include/gsrast.h)
        self.acc_clear_for = None

    def ensure(self, nbytes, keep_prefix=0):
        """Grow the scratch arena to >= nbytes, preserving its first keep_prefix bytes."""
        if self.scratch.numel() >= nbytes:
            return
        new = torch.empty(int(nbytes * 1.25) + 4096, dtype=torch.uint8, device=self.device)
        if keep_prefix:
            new[:keep_prefix].copy_(self.scratch[:keep_prefix])
        self.scratch = new
        self.regrows += 1
        self.acc_clear_for = None


_workspaces = {}
_MAX_WORKSPACES = 8        # arenas of streams that went away are dropped, least recently used first


def _workspace(device):
    # one arena per (device, stream): calls on different streams may run concurrently and must not
    # share scratch; calls on one stream are ordered
    index = device.index if device.index is not None else torch.cuda.current_device()
    key = (device.type, index, torch.cuda.current_stream(index).cuda_stream)
    ws = _workspaces.pop(key, None)
    if ws is None:
        ws = _DeviceWorkspace(device)
        while len(_workspaces) >= _MAX_WORKSPACES:
            # a short-lived stream must not pin an arena for the life of the process; the tensor is returned to
            # the caching allocator, which keeps it alive until the work queued on it has run
            _workspaces.pop(next(iter(_workspaces)))
    _workspaces[key] = ws                                # most recently used last
    return ws


def workspace_stats(device=None):
    """{'scratch_bytes', 'regrows', 'last_num_rendered'} of the current stream's arena on a device
    (config C5 reporting)."""
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    ws = _workspace(device)
    return dict(scratch_bytes=ws.scratch.numel(), regrows=ws.regrows, last_num_rendered=ws.last_R)


def _ptr(t):
    return None if t is None else t.data_ptr()


# View-parallel fused reduction (view_parallel.FusedGradientReducer): when set, the backward kernel
# adds its gradients into NVLink-multicast buffers instead of storing them locally.
_gradient_reducer = None


# Optional timeline probe (GSR_TRACE_EVENTS=1): CUDA events at the phase boundaries of a step on the current
# stream; trace_summary() turns them into mean gaps.  Diagnostic for the multi-GPU overhead, off by default.
_TRACE = os.environ.get("GSR_TRACE_EVENTS", "0") == "1"
_trace_events = []


def _trace(name, device):
    if _TRACE:
        ev = torch.cuda.Event(enable_timing=True)
        ev.record(torch.cuda.current_stream(device))
        _trace_events.append((name, ev))


def trace_reset():
    _trace_events.clear()


def trace_summary():
    """{'a->b': mean ms} over consecutive recorded events (call after a device synchronize)."""
    acc = {}
    for (n0, e0), (n1, e1) in zip(_trace_events, _trace_events[1:]):
        k = f"{n0}->{n1}"
        t, c = acc.get(k, (0.0, 0))
        acc[k] = (t + e0.elapsed_time(e1), c + 1)
    return {k: round(t / c, 4) for k, (t, c) in acc.items()}


def set_gradient_reducer(reducer):
    global _gradient_reducer
    _gradient_reducer = reducer


def _f32c(t, name):
    """fp32 contiguous CUDA tensor or None for an empty tensor (the reference's nullptr rule)."""
    if t is None or t.numel() == 0:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (this rasterizer has no CPU path)")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _params(P, M, rs, bg, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
            viewmatrix, projmatrix, campos, sh_dc=None, raw=0):
    p = GsrParams()
    p.P, p.D, p.M = P, int(rs.sh_degree), M
    p.raw_inputs = int(raw)
    p.W, p.H = int(rs.image_width), int(rs.image_height)
    p.tanfovx, p.tanfovy = float(rs.tanfovx), float(rs.tanfovy)
    p.scale_modifier = float(rs.scale_modifier)
    p.prefiltered, p.antialiasing, p.debug = int(bool(rs.prefiltered)), int(bool(rs.antialiasing)), int(bool(rs.debug))
    p.background, p.means3D, p.shs = _ptr(bg), _ptr(means3D), _ptr(sh)
    p.colors_precomp, p.opacities = _ptr(colors_precomp), _ptr(opacities)
    p.scales, p.rotations, p.cov3D_precomp = _ptr(scales), _ptr(rotations), _ptr(cov3Ds_precomp)
    p.viewmatrix, p.projmatrix, p.campos = _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos)
    p.sh_dc = _ptr(sh_dc)
    return p


def _forward_native(rs, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                    state_out=None, dc=None, raw=0, clear_accumulator=False):
    """Role of _C.rasterize_gaussians (ext.cpp:16 -> rasterize_points.cu:45-146).

    Returns (num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, invdepths).
    state_out: optional dict that receives views of the binning / image state (parity tests).
    dc: split SH storage — degree-0 coefficients (P,1,3); `sh` then holds coefficients 1..M-1.
    raw: bit mask _lib.RAW_*: scales / rotations / opacities are the model's raw parameters (fused activations).
    clear_accumulator: a backward will follow: the blend forward also zeroes the backward's accumulator."""
    lib = _lib.load()
    if means3D.ndim != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise RuntimeError("means3D must be a CUDA tensor (this rasterizer has no CPU path)")
    device = means3D.device
    P = means3D.shape[0]
    H, W = int(rs.image_height), int(rs.image_width)
    with torch.cuda.device(device):
        color = torch.empty((3, H, W), dtype=torch.float32, device=device)
        invdepth = torch.empty((1, H, W), dtype=torch.float32, device=device)
        radii = torch.empty((P,), dtype=torch.int32, device=device)
        empty = torch.empty(0, dtype=torch.uint8, device=device)
        if P == 0:                                   # rasterize_points.cu:107: no launch, zeros
            return 0, color.zero_(), radii, empty, empty, empty, invdepth.zero_()

        means3D = _f32c(means3D, "means3D")
        sh = _f32c(sh, "sh")
        colors_precomp = _f32c(colors_precomp, "colors_precomp")
        opacities = _f32c(opacities, "opacities")
        scales = _f32c(scales, "scales")
        rotations = _f32c(rotations, "rotations")
        cov3Ds_precomp = _f32c(cov3Ds_precomp, "cov3Ds_precomp")
        bg = _f32c(rs.bg, "bg")
        viewmatrix = _f32c(rs.viewmatrix, "viewmatrix")
        projmatrix = _f32c(rs.projmatrix, "projmatrix")
        campos = _f32c(rs.campos, "campos")
        dc = _f32c(dc, "dc")
        M = (sh.shape[1] if sh is not None else 0) + (1 if dc is not None else 0)
        p = _params(P, M, rs, bg, means3D, sh, colors_precomp, opacities, scales, rotations,
                    cov3Ds_precomp, viewmatrix, projmatrix, campos, dc, raw)

        ws = _workspace(device)
        stream = torch.cuda.current_stream(device)
        geom = torch.empty(lib.gsr_geom_bytes(P), dtype=torch.uint8, device=device)
        img = torch.empty(lib.gsr_img_bytes(W, H), dtype=torch.uint8, device=device)
        ws.ensure(lib.gsr_scratch_bytes(P, W, H, 0))        # no part of the arena depends on num_rendered

        buf = GsrBuffers()
        buf.geom, buf.geom_bytes = geom.data_ptr(), geom.numel()
        buf.img, buf.img_bytes = img.data_ptr(), img.numel()
        buf.scratch, buf.scratch_bytes = ws.scratch.data_ptr(), ws.scratch.numel()
        buf.num_rendered_host = ws.pinned.data_ptr()
        buf.num_rendered_event = ws.counted.cuda_event
        if ws.acc_clear_for != (P, W, H):
            ws.acc_clear_for = None          # this forward lays its buffers out differently: they may cover the zeros
        if clear_accumulator:
            buf.accumulator = _lib.ACC_CLEAR_IN_FORWARD
        out = GsrForwardOut(color.data_ptr(), invdepth.data_ptr(), radii.data_ptr())

        _lib.check(lib.gsr_forward_stage1(C.byref(p), C.byref(buf), C.byref(out), stream.cuda_stream))
        # The one host wait of the path (the reference blocks in cudaMemcpy, rasterizer_impl.cu:313).  The count
        # is ready as soon as the preprocess kernel has run: the depth sort and the tile counting go on while the
        # host sizes `bin` and enqueues stage 2.
        ws.counted.synchronize()
        R, R_hi = int(ws.pinned[0]), int(ws.pinned[1])
        if R_hi != 0 or R < 0:
            raise RuntimeError("num_rendered exceeds int32 (the reference stores it in an int)")
        ws.last_R = R
        binb = torch.empty(lib.gsr_bin_bytes(R), dtype=torch.uint8, device=device)
        buf.bin, buf.bin_bytes = binb.data_ptr(), binb.numel()
        _lib.check(lib.gsr_forward_stage2(C.byref(p), C.byref(buf), C.byref(out), R, stream.cuda_stream))
        if clear_accumulator:
            ws.acc_clear_for = (P, W, H)
        if state_out is not None:
            state_out.update(_inspect_state(lib, p, buf, R, geom, binb, img, stream))
    return R, color, radii, geom, binb, img, invdepth


def _inspect_state(lib, p, buf, R, geom, binb, img, stream):
    """Views of the forward state through the gsr_debug_* hooks: the reference's sorted key list
    (tile<<32 | depth bits), point_list, ranges, final_T, n_contrib and the 16-float records."""
    ptrs = [C.c_void_p() for _ in range(5)]
    _lib.check(lib.gsr_debug_state_ptrs(C.byref(p), C.byref(buf), R, *[C.byref(x) for x in ptrs]))

    def view(t, ptr, dtype, count):
        off = ptr.value - t.data_ptr()
        return t[off:off + count * torch.empty(0, dtype=dtype).element_size()].view(dtype)

    N = p.W * p.H
    T = ((p.W + 15) // 16) * ((p.H + 15) // 16)
    counts = torch.zeros(3, dtype=torch.int64, device=geom.device)
    if R > 0:
        _lib.check(lib.gsr_debug_count_work(C.byref(p), C.byref(buf), R, counts.data_ptr(), stream.cuda_stream))
    keys = torch.empty(max(R, 1), dtype=torch.int64, device=geom.device)
    if R > 0:
        _lib.check(lib.gsr_debug_sorted_keys(C.byref(p), C.byref(buf), R, keys.data_ptr(), stream.cuda_stream))
    return dict(point_list=view(binb, ptrs[0], torch.int32, R), ranges=view(img, ptrs[1], torch.int32, 2 * T).view(T, 2),
                final_T=view(img, ptrs[2], torch.float32, N), n_contrib=view(img, ptrs[3], torch.int32, N),
                records=view(geom, ptrs[4], torch.float32, 16 * p.P).view(p.P, 16), point_list_keys=keys[:R],
                work_counts=counts)


def _present(t):
    return t is not None and t.numel() > 0


def _grad_layout(P, M, colors_precomp, cov3Ds_precomp, sh, scales, rotations, split=False):
    """Names and shapes of the per-Gaussian gradient tensors the backward kernel writes.
    M counts all SH coefficients; split: coefficient 0 has its own tensor ("sh_dc"), "sh" holds the rest."""
    want = [("means3D", (P, 3), True), ("opacity", (P, 1), True),
            ("colors", (P, 3), _present(colors_precomp)), ("cov3D", (P, 6), _present(cov3Ds_precomp)),
            ("sh", (P, M - 1 if split else M, 3), _present(sh)), ("scales", (P, 3), _present(scales)),
            ("rotations", (P, 4), _present(rotations)), ("sh_dc", (P, 1, 3), split)]
    return [n for n, _, on in want if on], [shp for _, shp, on in want if on]


def _backward_native(rs, means3D, radii, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                     grad_out_color, grad_out_depth, sh, geom, R, binb, img, dc=None, raw=0):
    """Role of _C.rasterize_gaussians_backward (rasterize_points.cu:149-248).

    Returns (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales,
    dL_drotations); entries for absent inputs are None.  With split SH storage (`dc`), dL_dsh covers
    coefficients 1..M-1 and a ninth entry dL_ddc (P,1,3) is appended."""
    lib = _lib.load()
    device = means3D.device
    P = means3D.shape[0]
    H, W = int(rs.image_height), int(rs.image_width)
    f32 = dict(dtype=torch.float32, device=device)
    with torch.cuda.device(device):
        sh_c = _f32c(sh, "sh")
        dc_c = _f32c(dc, "dc")
        split = dc_c is not None
        M = (sh_c.shape[1] if sh_c is not None else 0) + (1 if split else 0)
        if P == 0:
            z = lambda *s: torch.zeros(*s, **f32)
            r0 = (z(0, 3), z(0, 3), z(0, 1), z(0, 3), z(0, 6), z(0, M - 1 if split else M, 3), z(0, 3), z(0, 4))
            return r0 + (z(0, 1, 3),) if split else r0
        means3D = _f32c(means3D, "means3D")
        colors_precomp = _f32c(colors_precomp, "colors_precomp")
        opacities = _f32c(opacities, "opacities")
        scales = _f32c(scales, "scales")
        rotations = _f32c(rotations, "rotations")
        cov3Ds_precomp = _f32c(cov3Ds_precomp, "cov3Ds_precomp")
        bg = _f32c(rs.bg, "bg")
        viewmatrix = _f32c(rs.viewmatrix, "viewmatrix")
        projmatrix = _f32c(rs.projmatrix, "projmatrix")
        campos = _f32c(rs.campos, "campos")
        grad_out_color = _f32c(grad_out_color, "grad_out_color")
        grad_out_depth = _f32c(grad_out_depth, "grad_out_depth")
        if grad_out_color is None:
            grad_out_color = torch.zeros((3, H, W), **f32)
        p = _params(P, M, rs, bg, means3D, sh_c, colors_precomp, opacities, scales, rotations,
                    cov3Ds_precomp, viewmatrix, projmatrix, campos, dc_c, raw)

        ws = _workspace(device)
        stream = torch.cuda.current_stream(device)
        ws.ensure(lib.gsr_scratch_bytes(P, W, H, R))
        buf = GsrBuffers()
        buf.geom, buf.geom_bytes = geom.data_ptr(), geom.numel()
        buf.img, buf.img_bytes = img.data_ptr(), img.numel()
        buf.bin, buf.bin_bytes = binb.data_ptr(), binb.numel()
        buf.scratch, buf.scratch_bytes = ws.scratch.data_ptr(), ws.scratch.numel()
        buf.num_rendered_host = ws.pinned.data_ptr()
        if ws.acc_clear_for == (P, W, H):    # zeroed by the forward of a view of this size, untouched since
            buf.accumulator = _lib.ACC_IS_CLEAR
        ws.acc_clear_for = None              # this backward dirties it

        dL_dmeans2D = torch.empty((P, 3), **f32)
        names, shapes = _grad_layout(P, M, colors_precomp, cov3Ds_precomp, sh_c, scales, rotations, split)
        red = _gradient_reducer
        exchange = red is not None and hasattr(red, "use_exchange") and red.use_exchange(names)
        if red is None or exchange:
            # every output is fully written (by the fused backward kernel, or by the expand pass): no zero fills
            outs = {n: torch.empty(shp, **f32) for n, shp in zip(names, shapes)}
            addr = {n: t.data_ptr() for n, t in outs.items()}
        else:
            # outputs are views of this rank's replica of a symmetric buffer; the kernel reduces into
            # all replicas through the multicast address
            views, addrs = red.buffers(shapes, device)
            outs = dict(zip(names, views))
            addr = dict(zip(names, addrs))
        if exchange:
            # factored exchange: the kernel publishes 64 B per contributing Gaussian through the multicast
            # address; after the rank barrier the expand pass rebuilds the dense, summed gradients locally
            local, mcast = red.exchange_buffers(P, device)
            _trace("bwd0", device)
            red.begin()
            _trace("begun", device)
            io = GsrBackwardIO(grad_out_color.data_ptr(), _ptr(grad_out_depth), None, dL_dmeans2D.data_ptr(),
                               None, None, None, None, None, None, int(getattr(red, "multicast", True)), 0,
                               mcast["records"], mcast["rgb"], mcast["campos"], None)
            _lib.check(lib.gsr_backward(C.byref(p), C.byref(buf), R, radii.data_ptr(), C.byref(io),
                                        stream.cuda_stream))
            _trace("sent", device)
            red.end()
            _trace("barrier", device)
            _lib.check(lib.gsr_exchange_expand(C.byref(p), red.world, local["records"], local["rgb"], local["campos"],
                                               addr["means3D"], addr["opacity"], addr["sh"], addr["scales"],
                                               addr["rotations"], addr.get("sh_dc"), stream.cuda_stream))
            _trace("expanded", device)
        else:
            if red is not None:
                red.begin()
            io = GsrBackwardIO(grad_out_color.data_ptr(), _ptr(grad_out_depth), addr["means3D"],
                               dL_dmeans2D.data_ptr(), addr["opacity"], addr.get("colors"), addr.get("cov3D"),
                               addr.get("sh"), addr.get("scales"), addr.get("rotations"),
                               0 if red is None else 1, 0 if red is None else int(getattr(red, "bands", 0)),
                               None, None, None, addr.get("sh_dc"))
            _lib.check(lib.gsr_backward(C.byref(p), C.byref(buf), R, radii.data_ptr(), C.byref(io),
                                        stream.cuda_stream))
            if red is not None:
                red.end()
        dL_dmeans3D, dL_dopacity = outs["means3D"], outs["opacity"]
        dL_dcolors, dL_dcov3D, dL_dsh = outs.get("colors"), outs.get("cov3D"), outs.get("sh")
        dL_dscales, dL_drotations = outs.get("scales"), outs.get("rotations")
    ret = (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)
    return ret + (outs["sh_dc"],) if split else ret


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings, dc=None, raw=0):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings, dc, raw)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings, dc=None, raw=0):
        """The reference's nine arguments (__init__.py:49-62) + `dc`: split SH storage (degree-0 coefficients
        (P,1,3) separate from `sh` = coefficients 1..M-1), read by the kernels in place — no concatenation;
        + `raw`: bit mask of the inputs that are the model's raw parameters (activations fused into the kernels,
        gradients returned w.r.t. the raw parameters)."""
        split = dc is not None
        _trace("fwd0", means3D.device)
        red = _gradient_reducer
        if red is not None and any(ctx.needs_input_grad) and means3D.shape[0] > 0 and hasattr(red, "prepare"):
            # view-parallel: clear the symmetric gradient buffer + rank barrier overlap the forward
            M = (sh.shape[1] if _present(sh) else 0) + (1 if split else 0)
            names, shapes = _grad_layout(means3D.shape[0], M, colors_precomp, cov3Ds_precomp, sh, scales, rotations, split)
            xchg = hasattr(red, "use_exchange") and red.use_exchange(names)
            red.prepare(shapes, means3D.device, exchange_P=means3D.shape[0] if xchg else None)
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, invdepths = _forward_native(
            raster_settings, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, dc=dc, raw=raw,
            clear_accumulator=any(ctx.needs_input_grad))
        ctx.raster_settings = raster_settings
        ctx.raw = raw
        ctx.num_rendered = num_rendered
        ctx.split = split
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities,
                              geomBuffer, binningBuffer, imgBuffer, *((dc,) if split else ()))
        ctx.mark_non_differentiable(radii)
        _trace("fwd1", means3D.device)
        return color, radii, invdepths

    @staticmethod
    def backward(ctx, grad_out_color, _, grad_out_depth):
        num_rendered = ctx.num_rendered
        raster_settings = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities,
         geomBuffer, binningBuffer, imgBuffer, *rest) = ctx.saved_tensors
        dc = rest[0] if ctx.split else None
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh,
         grad_scales, grad_rotations, *grad_dc) = _backward_native(
            raster_settings, means3D, radii, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
            grad_out_color, grad_out_depth, sh, geomBuffer, num_rendered, binningBuffer, imgBuffer, dc=dc, raw=ctx.raw)
        grads = (
            grad_means3D,
            grad_means2D,
            grad_sh,
            grad_colors_precomp,
            grad_opacities,
            grad_scales,
            grad_rotations,
            grad_cov3Ds_precomp,
            None,
            grad_dc[0] if grad_dc else None,
            None,
        )
        return grads


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        # Mark visible points (based on frustum culling for camera) with a boolean
        with torch.no_grad():
            raster_settings = self.raster_settings
            lib = _lib.load()
            if not positions.is_cuda:
                raise RuntimeError("positions must be a CUDA tensor (this rasterizer has no CPU path)")
            P = positions.shape[0]
            visible = torch.zeros((P,), dtype=torch.bool, device=positions.device)
            if P:
                with torch.cuda.device(positions.device):
                    pos = _f32c(positions, "positions")
                    view = _f32c(raster_settings.viewmatrix, "viewmatrix")
                    proj = _f32c(raster_settings.projmatrix, "projmatrix")
                    _lib.check(lib.gsr_mark_visible(P, pos.data_ptr(), view.data_ptr(), proj.data_ptr(),
                                                    visible.data_ptr(),
                                                    torch.cuda.current_stream(positions.device).cuda_stream))
        return visible

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, dc=None, raw_inputs=0):
        """raw_inputs (an extension, default off): bit mask RAW_SCALES | RAW_ROTATIONS | RAW_OPACITIES — the named
        inputs are the model's raw parameters (log-scales, unnormalised quaternions, opacity logits); the kernels
        apply exp / normalize / sigmoid themselves and return gradients w.r.t. the raw parameters.  Inputs that are
        `activations.Deferred` proxies (what activations.install(GaussianModel) makes get_scaling / get_rotation /
        get_opacity return) set their bit by themselves."""
        raster_settings = self.raster_settings
        from . import activations as _act
        opacities, scales, rotations, raw_inputs = _act.unwrap(opacities, scales, rotations, raw_inputs)

        # `dc=` is the keyword render() uses when SparseGaussianAdam is importable
        # (gaussian_renderer/__init__.py:115-125, separate_sh): degree-0 coefficients (P,1,3) and the rest
        # (P,15,3) arrive separately.  It is appended to the fork's signature so positional callers of the
        # fork's order keep working.
        if dc is not None:
            if colors_precomp is not None:
                raise Exception('Please provide exactly one of either SHs or precomputed colors!')
            if shs is None or shs.numel() == 0:
                shs, dc = dc, None               # degree-0 model: dc alone is the dense layout
            elif dc.dim() != 3 or dc.shape[1] != 1 or dc.shape[0] != shs.shape[0]:
                raise Exception('dc must have dimensions (num_points, 1, 3)')

        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide exactly one of either SHs or precomputed colors!')

        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])
        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])

        # Invoke C++/CUDA rasterization routine
        return rasterize_gaussians(
            means3D,
            means2D,
            shs,
            colors_precomp,
            opacities,
            scales,
            rotations,
            cov3D_precomp,
            raster_settings,
            dc,
            int(raw_inputs),
        )
