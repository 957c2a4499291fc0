"""View-parallel multi-GPU plumbing: one process per GPU, Gaussians replicated,
each rank rasterizes different cameras, per-Gaussian parameter gradients are summed with ONE
all-reduce over a flat fp32 buffer (NCCL over NVLink on GPUs; gloo in the CPU tests).

The reference has no distributed code at all; this is the exchange step north_star prescribes.
Densification statistics are reduced as per-view values because the reference sums per-view
norms (scene/gaussian_model.py:643-654) and takes a running max of radii (train.py:212).
"""
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist


def cameras_for_rank(num_views: int, rank: int, world_size: int) -> List[int]:
    """Rank r of G rasterizes cameras r, r+G, r+2G, ... of the batch."""
    return list(range(rank, num_views, world_size))


class FlatGradients:
    """Packs the gradients of a parameter list into one flat buffer and all-reduces it.

    After `all_reduce()` every parameter's .grad is a view into the flat buffer, so the
    optimizer reads the reduced values without an unpack pass.  The buffer is persistent and
    only reallocated when the total element count changes (densification)."""

    def __init__(self, params: Sequence[torch.Tensor], group: Optional[dist.ProcessGroup] = None):
        self.params = list(params)
        self.group = group
        self.flat = None
        self._views = None

    def _ensure(self):
        n = sum(p.numel() for p in self.params)
        p0 = self.params[0]
        if self.flat is None or self.flat.numel() != n or self.flat.device != p0.device:
            self.flat = torch.empty(n, dtype=torch.float32, device=p0.device)
            self._views = []
            off = 0
            for p in self.params:
                self._views.append(self.flat[off:off + p.numel()].view_as(p))
                off += p.numel()
        return self.flat

    def pack(self):
        self._ensure()
        srcs, dsts = [], []
        for p, v in zip(self.params, self._views):
            if p.grad is None:
                v.zero_()
            elif p.grad.data_ptr() != v.data_ptr():
                srcs.append(p.grad)
                dsts.append(v)
        if srcs:
            torch._foreach_copy_(dsts, srcs)
        return self.flat

    def all_reduce(self, average: bool = False, async_op: bool = False):
        """Sum (or mean) the packed gradients over the group; returns the work handle if async."""
        flat = self.pack()
        work = None
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=async_op)
            if average and not async_op:
                flat.div_(dist.get_world_size(self.group))
        for p, v in zip(self.params, self._views):
            p.grad = v
        return work

    def bytes_per_step(self) -> int:
        return 0 if self.flat is None else self.flat.numel() * 4


def reduce_densification_stats(viewspace_grad: torch.Tensor, radii: torch.Tensor,
                               group: Optional[dist.ProcessGroup] = None):
    """Per-view densification statistics summed / maxed over ranks.

    Returns (grad_norm_sum (P,1), denom (P,1), max_radii (P,)) matching
    add_densification_stats (scene/gaussian_model.py:643-654) applied once per view."""
    vis = radii > 0
    norm = torch.norm(viewspace_grad[:, :2], dim=-1, keepdim=True) * vis[:, None]
    stats = torch.cat([norm, vis[:, None].to(norm.dtype)], dim=1).contiguous()
    mx = radii.clone()
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
    return stats[:, 0:1], stats[:, 1:2], mx
