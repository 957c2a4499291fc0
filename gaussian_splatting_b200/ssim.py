"""Fused SSIM — host-side mirror of the reference's `fused_ssim` package
(submodules/fused-ssim/fused_ssim/__init__.py:1-41; used by train.py:34-38,155-159).

Same names and argument meaning: `FusedSSIMMap.apply(C1, C2, img1, img2, padding, train)` returns the
SSIM map, `fused_ssim(img1, img2, padding="same", train=True)` its mean.  Kernels: csrc/ssim.cu through
the C-ABI (gsr_ssim_forward / gsr_ssim_backward).  `fused_ssim` does not materialise the map: the
mean is reduced inside the forward launch and the backward spreads the scalar upstream gradient
in-kernel, which removes the map write, torch's mean kernel, the expand of its backward and the
`dL_dmap` read.  There is no CPU path.
"""
import torch

from . import _lib

allowed_padding = ["same", "valid"]
_CROP = {"same": 0, "valid": 5}


def _check(img1, img2):
    if not (img1.is_cuda and img2.is_cuda):
        raise RuntimeError("fused_ssim needs CUDA tensors (no CPU path)")
    if img1.dim() != 4 or img1.shape != img2.shape:
        raise RuntimeError(f"fused_ssim expects two [B, CH, H, W] images of equal shape, got "
                           f"{tuple(img1.shape)} and {tuple(img2.shape)}")
    if img1.dtype != torch.float32 or img2.dtype != torch.float32:
        raise RuntimeError("fused_ssim expects float32 images")       # the reference's .data<float>() throws too


def _params(C1, C2, img1, img2, crop):
    B, CH, H, W = img1.shape
    return _lib.GsrSsimParams(img1.data_ptr(), img2.data_ptr(), B, CH, H, W, float(C1), float(C2), crop)


def _stream(t):
    return torch.cuda.current_stream(t.device).cuda_stream


def _forward(C1, C2, img1, img2, crop, train, want_map, want_mean):
    """-> (map | None, mean | None, (dm_dmu1, dm_dsigma1_sq, dm_dsigma12) | None)"""
    lib = _lib.load()
    _check(img1, img2)
    img1, img2 = img1.contiguous(), img2.contiguous()
    B, CH, H, W = img1.shape
    smap = torch.empty_like(img1) if want_map else None
    dms = tuple(torch.empty_like(img1) for _ in range(3)) if train else None
    mean = partials = None
    if want_mean:
        mean = torch.empty((), device=img1.device, dtype=torch.float32)
        partials = torch.empty(max(1, lib.gsr_ssim_partials(B, CH, H, W)), device=img1.device, dtype=torch.float32)
    ptr = lambda t: None if t is None else t.data_ptr()  # noqa: E731
    with torch.cuda.device(img1.device):
        _lib.check(lib.gsr_ssim_forward(_params(C1, C2, img1, img2, crop), ptr(smap),
                                        *(map(ptr, dms) if train else (None, None, None)),
                                        ptr(mean), ptr(partials), _stream(img1)))
    return smap, mean, dms, img1, img2


def _backward(C1, C2, img1, img2, crop, dL_dmap, dL_dmean, dms):
    lib = _lib.load()
    if dms is None or len(dms) != 3 or dms[0].numel() != img1.numel():
        raise RuntimeError("fused_ssim backward needs a train=True forward (the derivative maps were not saved)")
    grad = torch.empty_like(img1)
    with torch.cuda.device(img1.device):
        _lib.check(lib.gsr_ssim_backward(_params(C1, C2, img1, img2, crop),
                                         None if dL_dmap is None else dL_dmap.data_ptr(),
                                         None if dL_dmean is None else dL_dmean.data_ptr(),
                                         dms[0].data_ptr(), dms[1].data_ptr(), dms[2].data_ptr(),
                                         grad.data_ptr(), _stream(img1)))
    return grad


class FusedSSIMMap(torch.autograd.Function):
    """The reference's autograd function (fused_ssim/__init__.py:8-33): returns the map."""

    @staticmethod
    def forward(ctx, C1, C2, img1, img2, padding="same", train=True):
        smap, _, dms, img1c, img2c = _forward(C1, C2, img1, img2, 0, train, True, False)
        if padding == "valid":
            smap = smap[:, :, 5:-5, 5:-5]
        ctx.save_for_backward(img1c.detach(), img2c, *(dms if train else ()))
        ctx.C1, ctx.C2, ctx.padding = C1, C2, padding
        return smap

    @staticmethod
    def backward(ctx, opt_grad):
        img1, img2, *dms = ctx.saved_tensors
        dL_dmap = opt_grad
        if ctx.padding == "valid":
            dL_dmap = torch.zeros_like(img1)
            dL_dmap[:, :, 5:-5, 5:-5] = opt_grad
        dL_dmap = dL_dmap.contiguous().float()
        grad = _backward(ctx.C1, ctx.C2, img1, img2, 0, dL_dmap, None, dms if dms else None)
        return None, None, grad, None, None, None


class _FusedSSIMMean(torch.autograd.Function):
    """mean(map[crop]) with the reduction and the gradient spreading inside the kernels."""

    @staticmethod
    def forward(ctx, C1, C2, img1, img2, padding, train):
        crop = _CROP[padding]
        _, mean, dms, img1c, img2c = _forward(C1, C2, img1, img2, crop, train, False, True)
        ctx.save_for_backward(img1c.detach(), img2c, *(dms if train else ()))
        ctx.C1, ctx.C2, ctx.crop = C1, C2, crop
        return mean

    @staticmethod
    def backward(ctx, opt_grad):
        img1, img2, *dms = ctx.saved_tensors
        g = opt_grad.to(device=img1.device, dtype=torch.float32).contiguous()
        grad = _backward(ctx.C1, ctx.C2, img1, img2, ctx.crop, None, g, dms if dms else None)
        return None, None, grad, None, None, None


def fused_ssim(img1, img2, padding="same", train=True):
    """fused_ssim/__init__.py:35-41."""
    C1 = 0.01 ** 2
    C2 = 0.03 ** 2
    assert padding in allowed_padding
    return _FusedSSIMMean.apply(C1, C2, img1, img2, padding, train)
