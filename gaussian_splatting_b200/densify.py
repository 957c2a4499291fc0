"""Adaptive density control on the device.

Host-side mirror of the density-control methods of the reference's `GaussianModel`
(scene/gaussian_model.py), same names, argument meaning and side effects, over libgsrast.so
(csrc/densify.cu).  Each function takes the model object first, so it works on the reference's own
`GaussianModel` instance (attributes `_xyz _features_dc _features_rest _opacity _scaling _rotation
optimizer xyz_gradient_accum denom max_radii2D percent_dense`, optimizer groups named
`xyz f_dc f_rest opacity scaling rotation`, gaussian_model.py:228-245):

    densify_and_prune(gm, max_grad, min_opacity, extent, max_screen_size, radii)   # :572-621
    prune_points(gm, mask)                                                        # :433-451
    add_densification_stats(gm, viewspace_point_tensor, update_filter, radii=None) # :630-633 (+ train.py:212)
    reset_opacity(gm)                                                             # :323-327
    install(GaussianModel)   # binds the four as methods, so the reference's train.py runs unmodified

The reference performs ~150 torch kernels per densification (boolean-mask gathers with a host
synchronisation each, torch.cat per tensor and per Adam moment, intermediate models for the clone, split
and prune steps).  Here: one plan (3 small kernels + ONE host wait for the four counts) and one apply
launch that moves all 6 parameter tensors and their 12 Adam moments straight to the final layout.

The one random draw of the path, `torch.normal(mean=0, std=stds)` (:521), is taken from `noise`
(2, >=P, 3) standard-normal numbers indexed [copy][source Gaussian]; by default it is drawn with
`torch.randn(..., generator=generator)` on the device.  View-parallel ranks pass the same seeded generator
(or broadcast `noise`) so that all replicas take identical decisions.

No CPU path: parameters must be CUDA tensors, libgsrast.so must be present.
"""
import ctypes as C

import torch
from torch import nn

from . import _lib

GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
_ATTR = {"xyz": "_xyz", "f_dc": "_features_dc", "f_rest": "_features_rest", "opacity": "_opacity",
         "scaling": "_scaling", "rotation": "_rotation"}
_KIND = {"xyz": _lib.DENSIFY_XYZ, "scaling": _lib.DENSIFY_SCALING}


def _require_cuda(t, what):
    if not t.is_cuda:
        raise RuntimeError(f"{what} must be a CUDA tensor (density control has no CPU path)")


_device_ctx = torch.cuda.device


def _stream(device):
    return torch.cuda.current_stream(device).cuda_stream


def _wait(device):
    torch.cuda.current_stream(device).synchronize()


def _pinned_counts():
    return torch.zeros(4, dtype=torch.int32).pin_memory()


def _f32(t, what):
    _require_cuda(t, what)
    if t.dtype != torch.float32:
        raise RuntimeError(f"{what} must be float32")
    return t if t.is_contiguous() else t.contiguous()


def _groups(gm):
    """{name: (group, param, state or None)} with the reference's one-parameter-per-group layout (:456)."""
    out = {}
    for group in gm.optimizer.param_groups:
        name = group.get("name")
        if name in GROUPS:
            assert len(group["params"]) == 1
            p = group["params"][0]
            out[name] = (group, p, gm.optimizer.state.get(p, None))
    missing = [g for g in GROUPS if g not in out]
    if missing:
        raise RuntimeError(f"optimizer has no parameter group named {missing}")
    return out


def _plan(lib, device, args, P):
    plan = torch.empty(int(lib.gsr_densify_plan_words(P)), dtype=torch.int32, device=device)
    counts = _pinned_counts()
    _lib.check(lib.gsr_densify_plan(C.byref(args), plan.data_ptr(), counts.data_ptr(), _stream(device)))
    _wait(device)                       # the one host wait: output shapes depend on the counts
    A, B, Cn, Pn = (int(x) for x in counts)
    if Pn < 0:
        raise RuntimeError("densification would exceed 2^31 - 1 Gaussians")
    return plan, A, B, Cn, Pn


def _apply(gm, lib, device, plan, P, A, B, Cn, Pn, groups, extra=(), noise=None, noise_stride=0):
    """Moves parameters, moments and `extra` [(src, kind)] tensors to the planned rows; rebinds the optimizer the
    way _prune_optimizer / cat_tensors_to_optimizer do (:414-431, :453-484). Returns the moved `extra` tensors."""
    args = _lib.GsrDensifyApplyArgs()
    args.P, args.A, args.B, args.C = P, A, B, Cn
    keep = []           # keeps sources alive until the launch is enqueued
    n = 0

    def add(src, kind):
        nonlocal n
        src = _f32(src.detach(), "model tensor")
        dst = torch.empty((Pn,) + tuple(src.shape[1:]), dtype=torch.float32, device=device)
        width = 1
        for d in src.shape[1:]:
            width *= int(d)
        if width == 0:                      # e.g. f_rest at SH degree 0: nothing to move
            return dst
        if n >= _lib.DENSIFY_MAX_TENSORS:
            raise RuntimeError("too many tensors for one gsr_densify_apply call")
        args.tensors[n].src, args.tensors[n].dst = src.data_ptr(), dst.data_ptr()
        args.tensors[n].width, args.tensors[n].kind = width, kind
        n += 1
        keep.append(src)
        return dst

    new = {}
    for name in GROUPS:
        _, p, state = groups[name]
        np_ = add(p, _KIND.get(name, _lib.DENSIFY_COPY))
        nm = nv = None
        if state is not None and "exp_avg" in state:
            nm = add(state["exp_avg"], _lib.DENSIFY_MOMENT)
            nv = add(state["exp_avg_sq"], _lib.DENSIFY_MOMENT)
        new[name] = (np_, nm, nv)
    moved = [add(src, kind) for src, kind in extra]
    args.n_tensors = n
    scal = _f32(groups["scaling"][1].detach(), "scaling")
    rot = _f32(groups["rotation"][1].detach(), "rotation")
    args.scaling, args.rotation = scal.data_ptr(), rot.data_ptr()
    if noise is not None:
        args.noise, args.noise_stride = noise.data_ptr(), noise_stride
    _lib.check(lib.gsr_densify_apply(C.byref(args), plan.data_ptr(), _stream(device)))
    # rebind: new nn.Parameter per group, the state dict object (with its `step`) moves to the new key
    for name in GROUPS:
        group, p, state = groups[name]
        np_, nm, nv = new[name]
        param = nn.Parameter(np_.requires_grad_(True))
        if state is not None:
            if nm is not None:
                state["exp_avg"], state["exp_avg_sq"] = nm, nv
            del gm.optimizer.state[p]
            gm.optimizer.state[param] = state
        group["params"][0] = param
        setattr(gm, _ATTR[name], param)
    return moved


@torch.no_grad()
def densify_and_prune(gm, max_grad, min_opacity, extent, max_screen_size, radii, noise=None, generator=None):
    """GaussianModel.densify_and_prune (:572-621): clone small / split large Gaussians whose mean view-space
    gradient reaches `max_grad`, then prune transparent and (if `max_screen_size`) oversized ones.  `radii` is
    accepted for signature parity: the reference only threads it through as tmp_radii and discards it (:613-614).
    Returns the row counts of the result {"kept", "clones", "split" (sources; 2 children each), "P"} for logging."""
    lib = _lib.load()
    groups = _groups(gm)
    xyz = groups["xyz"][1]
    _require_cuda(xyz, "GaussianModel._xyz")
    device = xyz.device
    P = int(xyz.shape[0])
    if P == 0:
        return {"kept": 0, "clones": 0, "split": 0, "P": 0}
    with _device_ctx(device):
        a = _lib.GsrDensifyPlanArgs()
        a.P = P
        a.use_screen_size = 1 if max_screen_size else 0
        a.grad_threshold = float(max_grad)
        a.dense_extent = float(gm.percent_dense * extent)
        a.min_opacity = float(min_opacity)
        a.max_screen_size = float(max_screen_size) if max_screen_size else 0.0
        a.world_extent_limit = float(0.1 * extent)
        accum = _f32(gm.xyz_gradient_accum, "xyz_gradient_accum")
        denom = _f32(gm.denom, "denom")
        scal = _f32(groups["scaling"][1].detach(), "scaling")
        opac = _f32(groups["opacity"][1].detach(), "opacity")
        a.xyz_gradient_accum, a.denom = accum.data_ptr(), denom.data_ptr()
        a.scaling, a.opacity = scal.data_ptr(), opac.data_ptr()
        plan, A, B, Cn, Pn = _plan(lib, device, a, P)
        stride = 0
        if Cn > 0:
            if noise is None:
                noise = torch.randn((2, P, 3), dtype=torch.float32, device=device, generator=generator)
            noise = _f32(noise, "noise")
            if noise.dim() != 3 or noise.shape[0] != 2 or noise.shape[1] < P or noise.shape[2] != 3:
                raise RuntimeError("noise must have shape (2, >=P, 3)")
            stride = int(noise.shape[1])
        else:
            noise = None
        _apply(gm, lib, device, plan, P, A, B, Cn, Pn, groups, noise=noise, noise_stride=stride)
        # densification_postfix resets all three statistics (:506-508); pruning keeps them zero
        gm.xyz_gradient_accum = torch.zeros((Pn, 1), device=device)
        gm.denom = torch.zeros((Pn, 1), device=device)
        gm.max_radii2D = torch.zeros((Pn,), device=device)
        gm.tmp_radii = None
    return {"kept": A, "clones": B, "split": Cn, "P": Pn}


@torch.no_grad()
def prune_points(gm, mask):
    """GaussianModel.prune_points (:433-451): drop the rows where `mask` is true from the parameters, both Adam
    moments and the three statistics."""
    lib = _lib.load()
    groups = _groups(gm)
    xyz = groups["xyz"][1]
    _require_cuda(xyz, "GaussianModel._xyz")
    device = xyz.device
    P = int(xyz.shape[0])
    if P == 0:
        return
    with _device_ctx(device):
        _require_cuda(mask, "mask")
        m = mask if mask.dtype in (torch.bool, torch.uint8) else mask != 0
        m = m.reshape(-1).contiguous()
        if m.numel() != P:
            raise RuntimeError("mask must have one entry per Gaussian")
        a = _lib.GsrDensifyPlanArgs()
        a.P = P
        a.prune_mask = m.data_ptr()
        plan, A, B, Cn, Pn = _plan(lib, device, a, P)
        extra = [(gm.xyz_gradient_accum, _lib.DENSIFY_COPY), (gm.denom, _lib.DENSIFY_COPY),
                 (gm.max_radii2D, _lib.DENSIFY_COPY)]
        tmp = getattr(gm, "tmp_radii", None)
        moved = _apply(gm, lib, device, plan, P, A, B, Cn, Pn, groups, extra=extra)
        gm.xyz_gradient_accum, gm.denom, gm.max_radii2D = moved
        if tmp is not None:
            gm.tmp_radii = tmp[~m]


@torch.no_grad()
def add_densification_stats(gm, viewspace_point_tensor, update_filter, radii=None):
    """GaussianModel.add_densification_stats (:630-633); with `radii` also train.py:212-213
    (`max_radii2D[vis] = max(max_radii2D[vis], radii[vis])`) in the same launch.
    `update_filter` is what render() returns as visibility_filter — the int64 index tensor `(radii > 0).nonzero()`
    (gaussian_renderer/__init__.py:152) — or a boolean mask, or None with `radii` (then radii > 0 is the filter)."""
    lib = _lib.load()
    grad = viewspace_point_tensor.grad
    if grad is None:
        raise RuntimeError("viewspace_point_tensor has no gradient (call backward first)")
    grad = _f32(grad, "viewspace_point_tensor.grad")
    P = int(grad.shape[0])
    if P == 0:
        return
    if grad.dim() != 2 or grad.shape[1] != 3:
        raise RuntimeError("viewspace_point_tensor.grad must have shape (P, 3)")
    device = grad.device
    vis_ptr = idx_ptr = None
    n_idx = 0
    keep = None
    if update_filter is None:
        if radii is None:
            raise RuntimeError("add_densification_stats needs update_filter or radii")
    elif update_filter.dtype in (torch.bool, torch.uint8):
        keep = update_filter.reshape(-1).contiguous()
        if keep.numel() != P:
            raise RuntimeError("a boolean update_filter must have one entry per Gaussian")
        vis_ptr = keep.data_ptr()
    else:
        keep = update_filter.reshape(-1).to(torch.int64).contiguous()
        n_idx = int(keep.numel())
        if n_idx == 0:
            return
        idx_ptr = keep.data_ptr()
    if keep is not None:
        _require_cuda(keep, "update_filter")
    accum = gm.xyz_gradient_accum
    den = gm.denom
    for t, what in ((accum, "xyz_gradient_accum"), (den, "denom")):
        _require_cuda(t, what)
        if t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != P:
            raise RuntimeError(f"{what} must be a contiguous float32 tensor with one entry per Gaussian")
    rad_ptr = maxr_ptr = None
    if radii is not None:
        _require_cuda(radii, "radii")
        radii = radii if radii.dtype == torch.int32 else radii.to(torch.int32)
        radii = radii.reshape(-1).contiguous()
        maxr = gm.max_radii2D
        if radii.numel() != P or maxr.dtype != torch.float32 or not maxr.is_contiguous() or maxr.numel() != P:
            raise RuntimeError("radii / max_radii2D must have one entry per Gaussian (int32 / contiguous float32)")
        rad_ptr, maxr_ptr = radii.data_ptr(), maxr.data_ptr()
    with _device_ctx(device):
        _lib.check(lib.gsr_densify_stats(P, grad.data_ptr(), vis_ptr, idx_ptr, n_idx, rad_ptr, accum.data_ptr(),
                                         den.data_ptr(), maxr_ptr, _stream(device)))


@torch.no_grad()
def reset_opacity(gm):
    """GaussianModel.reset_opacity (:323-327) with replace_tensor_to_optimizer (:400-412): opacities are capped at
    0.01 (in activation space) and both Adam moments of the group are zeroed; the parameter object is replaced."""
    lib = _lib.load()
    groups = _groups(gm)
    group, p, state = groups["opacity"]
    _require_cuda(p, "GaussianModel._opacity")
    device = p.device
    new = _f32(p.detach(), "opacity").clone()
    P = int(new.shape[0])
    m = v = None
    if state is not None and "exp_avg" in state:
        m, v = torch.empty_like(new), torch.empty_like(new)
    with _device_ctx(device):
        _lib.check(lib.gsr_reset_opacity(P, new.data_ptr(), None if m is None else m.data_ptr(),
                                         None if v is None else v.data_ptr(), _stream(device)))
    param = nn.Parameter(new.requires_grad_(True))
    if state is not None:
        if m is not None:
            state["exp_avg"], state["exp_avg_sq"] = m, v
        del gm.optimizer.state[p]
        gm.optimizer.state[param] = state
    group["params"][0] = param
    gm._opacity = param


def install(model_class):
    """Binds the device implementations as methods of the reference's GaussianModel class:
        from scene.gaussian_model import GaussianModel
        import gaussian_splatting_b200.densify as d; d.install(GaussianModel)
    after which train.py:210-230 runs unmodified on the fused kernels."""
    model_class.densify_and_prune = densify_and_prune
    model_class.prune_points = prune_points
    model_class.add_densification_stats = add_densification_stats
    model_class.reset_opacity = reset_opacity
    return model_class
