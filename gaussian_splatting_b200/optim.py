"""SparseGaussianAdam — the optimizer the fork probes for (train.py:41-45,
scene/gaussian_model.py:25-28) and constructs as `SparseGaussianAdam(l, lr=0.0, eps=1e-15)`
(gaussian_model.py:246-251), stepped with `optimizer.step(visible, N)` (train.py:240-247), but
does not ship.

One fused kernel per parameter tensor (libgsrast.so: gsr_adam_update, csrc/adam.cu) instead of
torch's multi-pass foreach Adam.  Update rule = torch.optim.Adam (betas from the group, bias
correction with a per-parameter step count, no weight decay / amsgrad / maximize) on the rows of
visible Gaussians; invisible rows (whose gradient is zero) keep parameter and moments.  State
layout (`step`, `exp_avg`, `exp_avg_sq`) is torch.optim.Adam's, so the densification code that
edits optimizer state (gaussian_model.py:400-470) works unchanged.
"""
import math

import torch

from . import _lib


class SparseGaussianAdam(torch.optim.Adam):
    def __init__(self, params, lr, eps):
        super().__init__(params=params, lr=lr, eps=eps)

    @torch.no_grad()
    def step(self, visibility=None, N=None):
        lib = _lib.load()
        for group in self.param_groups:
            lr, eps = float(group["lr"]), float(group["eps"])
            beta1, beta2 = group["betas"]
            for param in group["params"]:
                if param.grad is None:
                    continue
                if not param.is_cuda:
                    raise RuntimeError("SparseGaussianAdam needs CUDA parameters (no CPU path)")
                if param.dtype != torch.float32 or not param.is_contiguous():
                    raise RuntimeError("SparseGaussianAdam needs contiguous float32 parameters")
                state = self.state[param]
                if len(state) == 0:
                    state["step"] = torch.tensor(0.0, dtype=torch.float32)
                    state["exp_avg"] = torch.zeros_like(param, memory_format=torch.preserve_format)
                    state["exp_avg_sq"] = torch.zeros_like(param, memory_format=torch.preserve_format)
                state["step"] += 1
                t = float(state["step"])
                n = param.shape[0] if N is None else int(N)
                if n == 0:
                    continue
                M = param.numel() // n
                grad = param.grad if param.grad.is_contiguous() else param.grad.contiguous()
                vis = None
                if visibility is not None:
                    vis = visibility if visibility.dtype in (torch.bool, torch.uint8) else visibility != 0
                    vis = vis.contiguous()
                with torch.cuda.device(param.device):
                    _lib.check(lib.gsr_adam_update(
                        param.data_ptr(), grad.data_ptr(), state["exp_avg"].data_ptr(), state["exp_avg_sq"].data_ptr(),
                        None if vis is None else vis.data_ptr(), lr, float(beta1), float(beta2), eps,
                        1.0 - beta1 ** t, math.sqrt(1.0 - beta2 ** t), n, M,
                        torch.cuda.current_stream(param.device).cuda_stream))
