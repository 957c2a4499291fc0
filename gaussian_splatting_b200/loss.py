"""The training loss of the reference in one forward and one backward launch (csrc/ssim.cu: loss_fwd_kernel /
loss_bwd_kernel through gsr_loss_forward / gsr_loss_backward).

Replaces, for the image the rasterizer just produced, `l1_loss` (utils/loss_utils.py:40), `fused_ssim`
(train.py:155-159), the weighted sum of train.py:162 and the inverse-depth regulariser of train.py:165-174 — in the
reference about ten elementwise / reduction kernels forward and as many backward around the two fused-ssim launches:

    Ll1 = l1_loss(image, gt_image)
    loss = (1.0 - opt.lambda_dssim) * Ll1 + opt.lambda_dssim * (1.0 - fused_ssim(image[None], gt_image[None]))
    loss += depth_l1_weight(iteration) * torch.abs((invDepth - mono_invdepth) * depth_mask).mean()

becomes

    loss, parts = training_loss(image, gt_image, opt.lambda_dssim, invdepth=invDepth, mono_invdepth=mono_invdepth,
                                depth_mask=depth_mask, depth_weight=depth_l1_weight(iteration), return_parts=True)
    # parts = [Ll1, ssim_value, Ll1depth_pure] (detached, on the device)

There is no CPU path.
"""
import torch

from . import _lib


def _c(t, name):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"training_loss: {name} must be a CUDA tensor (no CPU path)")
    if t.dtype != torch.float32:
        raise RuntimeError(f"training_loss: {name} must be float32")
    return t.contiguous()


class _TrainingLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, invdepth, gt, mono, mask, lambda_dssim, depth_weight):
        lib = _lib.load()
        image, gt, invdepth, mono, mask = (_c(image, "image"), _c(gt, "gt"), _c(invdepth, "invdepth"),
                                           _c(mono, "mono_invdepth"), _c(mask, "depth_mask"))
        if image.dim() != 3 or image.shape != gt.shape:
            raise RuntimeError(f"training_loss expects image and gt of equal shape [CH, H, W], got {tuple(image.shape)} "
                               f"and {tuple(gt.shape)}")
        CH, H, W = image.shape
        for t, n in ((invdepth, "invdepth"), (mono, "mono_invdepth"), (mask, "depth_mask")):
            if t is not None and t.numel() != H * W:
                raise RuntimeError(f"training_loss: {n} must hold H*W = {H * W} values, got {tuple(t.shape)}")
        dev = image.device
        dms = torch.empty((3, CH, H, W), device=dev, dtype=torch.float32)
        out = torch.empty(4, device=dev, dtype=torch.float32)
        part = torch.empty(max(1, lib.gsr_loss_partials(CH, H, W)), device=dev, dtype=torch.float32)
        ptr = lambda t: None if t is None else t.data_ptr()  # noqa: E731
        p = _lib.GsrLossParams(image.data_ptr(), gt.data_ptr(), ptr(invdepth), ptr(mono), ptr(mask), CH, H, W,
                               float(lambda_dssim), float(depth_weight), 0.01 ** 2, 0.03 ** 2, 0)
        with torch.cuda.device(dev):
            _lib.check(lib.gsr_loss_forward(p, dms[0].data_ptr(), dms[1].data_ptr(), dms[2].data_ptr(), out.data_ptr(),
                                            part.data_ptr(), torch.cuda.current_stream(dev).cuda_stream))
        ctx.save_for_backward(image.detach(), gt, dms, *(t for t in (invdepth, mono, mask) if t is not None))
        ctx.have = (invdepth is not None, mask is not None)
        ctx.coef = (float(lambda_dssim), float(depth_weight))
        ctx.depth_shape = None if invdepth is None else invdepth.shape
        parts = out[1:]
        ctx.mark_non_differentiable(parts)
        return out[0], parts

    @staticmethod
    def backward(ctx, g, _parts):
        lib = _lib.load()
        image, gt, dms, *rest = ctx.saved_tensors
        have_depth, have_mask = ctx.have
        invdepth = rest[0].detach() if have_depth else None
        mono = rest[1] if have_depth else None
        mask = rest[2] if have_mask else None
        CH, H, W = image.shape
        dev = image.device
        g = g.to(device=dev, dtype=torch.float32).contiguous()
        d_image = torch.empty_like(image)
        d_depth = torch.empty(ctx.depth_shape, device=dev, dtype=torch.float32) if have_depth else None
        ptr = lambda t: None if t is None else t.data_ptr()  # noqa: E731
        p = _lib.GsrLossParams(image.data_ptr(), gt.data_ptr(), ptr(invdepth), ptr(mono), ptr(mask), CH, H, W,
                               ctx.coef[0], ctx.coef[1], 0.01 ** 2, 0.03 ** 2, 0)
        with torch.cuda.device(dev):
            _lib.check(lib.gsr_loss_backward(p, g.data_ptr(), dms[0].data_ptr(), dms[1].data_ptr(), dms[2].data_ptr(),
                                             d_image.data_ptr(), ptr(d_depth), torch.cuda.current_stream(dev).cuda_stream))
        return d_image, d_depth, None, None, None, None, None


def training_loss(image, gt, lambda_dssim, invdepth=None, mono_invdepth=None, depth_mask=None, depth_weight=0.0,
                  return_parts=False):
    """(1 - lambda) * L1 + lambda * (1 - SSIM) [+ depth_weight * inverse-depth L1] of train.py:155-174 as one scalar
    tensor that backpropagates to `image` and `invdepth`.  return_parts: also [L1, mean SSIM, depth L1] (detached)."""
    if (invdepth is None) != (mono_invdepth is None):
        raise RuntimeError("training_loss: invdepth and mono_invdepth come together")
    loss, parts = _TrainingLoss.apply(image, invdepth, gt, mono_invdepth, depth_mask, lambda_dssim, depth_weight)
    return (loss, parts) if return_parts else loss
