"""View-parallel multi-GPU plumbing: one process per GPU, Gaussians replicated,
each rank rasterizes different cameras, per-Gaussian parameter gradients are summed with ONE
all-reduce over a flat fp32 buffer (NCCL over NVLink on GPUs; gloo in the CPU tests).

The reference has no distributed code at all; this is the exchange step north_star prescribes.
Densification statistics are reduced as per-view values because the reference sums per-view
norms (scene/gaussian_model.py:643-654) and takes a running max of radii (train.py:212).
"""
import os
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
        if average and async_op:
            raise ValueError("FlatGradients.all_reduce: average=True needs the result (async_op=False); divide after wait()")
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


def accumulate_densification_stats(gm, viewspace_grad: torch.Tensor, radii: torch.Tensor,
                                   group: Optional[dist.ProcessGroup] = None):
    """train.py:212-215 for all views of a view-parallel step at once: every rank adds the statistics of ALL ranks'
    cameras to its replica's accumulators, so the replicas stay identical."""
    nsum, den, mx = reduce_densification_stats(viewspace_grad, radii, group)
    gm.xyz_gradient_accum += nsum
    gm.denom += den
    gm.max_radii2D = torch.max(gm.max_radii2D, mx.to(gm.max_radii2D.dtype))


def replica_generator(device, seed: int, iteration: int) -> torch.Generator:
    """The generator every rank seeds identically for the split offsets of one densification
    (the reference draws them with the global CUDA generator, gaussian_model.py:521)."""
    g = torch.Generator(device=device)
    g.manual_seed((int(seed) * 1000003 + int(iteration)) & 0x7FFFFFFFFFFFFFFF)
    return g


def densify_and_prune_replicated(gm, max_grad, min_opacity, extent, max_screen_size, radii, seed: int, iteration: int,
                                 group: Optional[dist.ProcessGroup] = None, verify: bool = True):
    """GaussianModel.densify_and_prune (gaussian_model.py:572-621) on every replica of a view-parallel job.
    No exchange is needed: parameters, optimizer state and (reduced) statistics are identical on all ranks and the
    one random draw comes from `replica_generator`, so every rank takes the same decisions and ends with bit-identical
    tensors.  `verify` cross-checks the four row counts over the ranks (one 32-byte all-reduce) and raises if a
    replica diverged (e.g. statistics accumulated without reduce_densification_stats)."""
    from . import densify
    device = gm._xyz.device
    info = densify.densify_and_prune(gm, max_grad, min_opacity, extent, max_screen_size, radii,
                                     generator=replica_generator(device, seed, iteration))
    if verify and dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        c = torch.tensor([info["kept"], info["clones"], info["split"], info["P"]], dtype=torch.int64, device=device)
        both = torch.cat((c, -c))
        dist.all_reduce(both, op=dist.ReduceOp.MAX, group=group)     # max(c) and -min(c) in one collective
        if not torch.equal(both[:4], -both[4:]):
            raise RuntimeError(f"view-parallel replicas diverged in densify_and_prune: rank counts {c.tolist()}, "
                               f"max {both[:4].tolist()}, min {(-both[4:]).tolist()}")
    return info


class FusedGradientReducer:
    """All-reduce of the rasterizer's input gradients INSIDE the backward kernel (NVLS multicast).

    The fused per-Gaussian backward kernel (csrc/preprocess_bwd.cu) writes every gradient exactly
    once.  With this reducer installed (`install()`), those writes become `multimem.red.add.f32` on
    the NVLink multicast address of a symmetric buffer replicated on all ranks, so the NVSwitch adds
    the ranks' contributions while the kernel is still computing; no separate NCCL all-reduce pass
    and no pack copy.  The gradients the autograd Function returns are views of this rank's replica
    and already hold the SUM over ranks.  Because the activation / concat backward that follows in
    autograd is linear in the incoming gradient and the parameters are replicated, reducing here is
    equivalent to reducing the leaf gradients (up to fp32 rounding).

    Contract (same as any gradient bucket that is reused): every rank calls backward the same number
    of times, and the gradients returned by one backward are consumed (optimizer step / copy /
    `grad = None`) before the next differentiable forward on this device (the replica is cleared
    during that forward).  dL_dmeans2D (per-view densification
    statistic) is never reduced.

    PyTorch is plumbing here: symmetric allocation + rendezvous + stream barriers come from
    torch.distributed._symmetric_memory; the reduction itself is in our kernel.
    """

    ALIGN = 32   # floats (128 B) between segments

    def __init__(self, group: Optional[dist.ProcessGroup] = None, bands: int = 0):
        """bands: tile-row bands of the blend backward pipelined against the NVLink-bound per-Gaussian
        kernel (0 = library default, 1 = no pipelining)."""
        import torch.distributed._symmetric_memory as symm_mem
        self.bands = int(os.environ.get("GSR_REDUCE_BANDS", bands))
        if self.bands <= 0:
            # measured on 8 x B200 (profiles/r01_scaling): 4 bands 4.19 ms/step vs 4.40 unbanded at 8 ranks;
            # at 2 ranks NVLink is not the bound and banding only adds contention (3.92 vs 3.69 ms)
            self.bands = 4 if dist.get_world_size(group if group is not None else dist.group.WORLD) >= 8 else 1
        self._symm_mem = symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.flat = None
        self.hdl = None
        self.key = None
        self.bytes_reduced = 0
        # GSR_REDUCE_EARLY=0: zero + barrier at the start of the backward instead of during the forward
        self.early = os.environ.get("GSR_REDUCE_EARLY", "1") != "0"
        # GSR_REDUCE_FACTORED=0: all-reduce the full SH gradients (236 B / Gaussian) instead of the factored
        # exchange (64 B / Gaussian + a local expand pass)
        self.factored = os.environ.get("GSR_REDUCE_FACTORED", "1") != "0"
        self._side = self._ready = None
        self._prepared = False
        self.self_cleaning = False

    @staticmethod
    def available() -> bool:
        try:
            import torch.distributed._symmetric_memory  # noqa: F401
        except Exception:   # noqa: BLE001
            return False
        return dist.is_available() and dist.is_initialized() and torch.cuda.is_available()

    def _layout(self, shapes):
        offs, off = [], 0
        for shp in shapes:
            n = 1
            for d in shp:
                n *= d
            offs.append(off)
            off += (n + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        return offs, off

    def buffers(self, shapes, device):
        """Returns (local views, multicast addresses) for gradient tensors of the given shapes.
        (Re)allocates + rendezvous (a collective) when the shapes change, e.g. after densification."""
        key = (tuple(tuple(s) for s in shapes), str(device))
        if key != self.key:
            offs, total = self._layout(shapes)
            self.flat = self._symm_mem.empty(max(total, self.ALIGN), dtype=torch.float32, device=device)
            self.hdl = self._symm_mem.rendezvous(self.flat, self.group.group_name)
            if not self.hdl.multicast_ptr:
                raise RuntimeError("NVLink multicast (NVLS) is not available for this group: use FlatGradients")
            self.key, self.offs, self.total = key, offs, total
            self.self_cleaning = False
        views, addrs = [], []
        for shp, off in zip(shapes, self.offs):
            n = 1
            for d in shp:
                n *= d
            views.append(self.flat[off:off + n].view(*shp))
            addrs.append(self.hdl.multicast_ptr + 4 * off)
        return views, addrs

    EXCHANGE_NAMES = ["means3D", "opacity", "sh", "scales", "rotations"]

    def use_exchange(self, names) -> bool:
        """Factored exchange (csrc/exchange.cu): SH + scale/rotation inputs, i.e. the training configuration."""
        return self.factored and [n for n in names if n != "sh_dc"] == self.EXCHANGE_NAMES

    def exchange_buffers(self, P, device):
        """Symmetric buffer of the factored exchange: records [P][12] | rgb [world][P][4] | campos [world][4].
        Returns (local addresses, multicast addresses) as dicts with keys records / rgb / campos; the
        multicast rgb / campos addresses point at THIS rank's slot.  (Re)allocates + rendezvous (a
        collective) when P changes, e.g. after densification."""
        key = ("xchg", int(P), str(device))
        if key != self.key:
            al = lambda n: (n + self.ALIGN - 1) // self.ALIGN * self.ALIGN  # noqa: E731
            o_rgb = al(12 * P)
            o_cam = o_rgb + al(4 * P * self.world)
            total = o_cam + al(4 * self.world)
            self.flat = self._symm_mem.empty(total, dtype=torch.float32, device=device)
            self.hdl = self._symm_mem.rendezvous(self.flat, self.group.group_name)
            if not self.hdl.multicast_ptr:
                raise RuntimeError("NVLink multicast (NVLS) is not available for this group: use FlatGradients")
            self.key, self.offs, self.total = key, [0, o_rgb, o_cam], total
            # zeroed ONCE: afterwards the expand pass resets exactly the rows that received something
            self.flat.zero_()
            self.self_cleaning = True
        rank = dist.get_rank(self.group)
        base, mc = self.flat.data_ptr(), self.hdl.multicast_ptr
        o_rec, o_rgb, o_cam = self.offs
        local = dict(records=base + 4 * o_rec, rgb=base + 4 * o_rgb, campos=base + 4 * o_cam)
        mcast = dict(records=mc + 4 * o_rec, rgb=mc + 4 * (o_rgb + 4 * P * rank), campos=mc + 4 * (o_cam + 4 * rank))
        return local, mcast

    def prepare(self, shapes, device, exchange_P=None):
        """Called from the rasterizer's forward (grad mode): zero this rank's replica and run the
        all-ranks-zeroed barrier on a side stream, overlapped with the forward pass, so the backward
        starts reducing at once.  Ordered after everything already enqueued on the current stream,
        i.e. after the consumers of the previous step's gradients."""
        if not self.early:
            return
        if exchange_P is not None:
            self.exchange_buffers(exchange_P, device)
        else:
            self.buffers(shapes, device)
        main = torch.cuda.current_stream(device)
        if self._side is None:
            self._side = torch.cuda.Stream(device=device)
            self._ready = torch.cuda.Event()
        self._side.wait_stream(main)
        with torch.cuda.stream(self._side):
            if not self.self_cleaning:
                self.flat.zero_()
            self.hdl.barrier(channel=0)
            self._ready.record(self._side)
        self._prepared = True

    def begin(self):
        """Zero this rank's replica and wait until every rank has done so (on the current stream);
        already done during the forward when prepare() ran."""
        if self._prepared:
            torch.cuda.current_stream(self.flat.device).wait_event(self._ready)
            self._prepared = False
            return
        if not self.self_cleaning:
            self.flat.zero_()
        self.hdl.barrier(channel=0)

    def end(self):
        """Wait until every rank's reductions have landed (on the current stream)."""
        self.hdl.barrier(channel=1)
        self.bytes_reduced = self.total * 4

    def install(self):
        from . import rasterizer
        rasterizer.set_gradient_reducer(self)
        return self

    def uninstall(self):
        from . import rasterizer
        rasterizer.set_gradient_reducer(None)
