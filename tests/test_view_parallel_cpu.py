"""World-size-2 gloo tests of the view-parallel exchange step on CPU."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gaussian_splatting_b200.view_parallel import FlatGradients, cameras_for_rank, reduce_densification_stats


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(3)
        P = 37
        shapes = [(P, 3), (P, 16, 3), (P, 1), (P, 3), (P, 4)]
        params = [torch.nn.Parameter(torch.randn(*s, generator=g)) for s in shapes]
        per_rank = []
        for r in range(world):
            gr = torch.Generator().manual_seed(100 + r)
            per_rank.append([torch.randn(*s, generator=gr) for s in shapes])
        for p, gt in zip(params, per_rank[rank]):
            p.grad = gt.clone()
        params[3].grad = None                                # a parameter with no gradient this step
        fg = FlatGradients(params)
        fg.all_reduce()
        ok = True
        for i, p in enumerate(params):
            want = sum(per_rank[r][i] for r in range(world)) if i != 3 else torch.zeros(shapes[3])
            ok &= torch.allclose(p.grad, want, atol=1e-6)
            ok &= p.grad.data_ptr() >= fg.flat.data_ptr()    # grads alias the flat buffer
        ok &= fg.bytes_per_step() == sum(p.numel() for p in params) * 4

        vg = torch.zeros(P, 3)
        vg[:, 0] = rank + 1.0
        radii = torch.full((P,), rank + 1, dtype=torch.int32)
        radii[rank::world] = 0                               # a different subset invisible per rank
        nsum, den, mx = reduce_densification_stats(vg, radii)
        want_den = torch.zeros(P, 1)
        want_n = torch.zeros(P, 1)
        for r in range(world):
            vis = torch.ones(P, dtype=torch.bool)
            vis[r::world] = False
            want_den[:, 0] += vis
            want_n[:, 0] += vis * (r + 1.0)
        ok &= torch.allclose(den, want_den) and torch.allclose(nsum, want_n)
        ok &= int(mx.max()) == world
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_flat_gradient_allreduce_world2():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert dict(out) == {0: True, 1: True}


def _densify_worker(rank, world, port, out, diverge):
    """Replicated density control on 2 ranks (kernels served by the CPU emulation, as in test_densify_emul_cpu.py)."""
    import contextlib
    from gaussian_splatting_b200 import _lib, densify
    from gaussian_splatting_b200.view_parallel import accumulate_densification_stats, densify_and_prune_replicated
    from tests import test_densify_emul_cpu as E
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        fake = E.FakeLib(E.load_emul())
        _lib.load = lambda: fake
        densify._require_cuda = lambda t, what: None
        densify._device_ctx = lambda d: contextlib.nullcontext()
        densify._stream = lambda d: None
        densify._wait = lambda d: None
        densify._pinned_counts = lambda: torch.zeros(4, dtype=torch.int32)
        gm = E.make_model("a_in_")                         # identical replicas
        P = gm._xyz.shape[0]
        gm.xyz_gradient_accum.zero_(), gm.denom.zero_(), gm.max_radii2D.zero_()
        # each rank saw a different camera: per-view gradients and radii differ, the reduced statistics do not
        g = torch.Generator().manual_seed(50 + rank)
        radii = (torch.rand(P, generator=g) * 30).to(torch.int32) * (torch.rand(P, generator=g) < 0.8)
        vgrad = torch.randn(P, 3, generator=g) * 0.0006
        accumulate_densification_stats(gm, vgrad, radii)
        if diverge and rank == 1:
            gm.xyz_gradient_accum += 1.0                   # a replica whose statistics were not reduced
        try:
            info = densify_and_prune_replicated(gm, 0.0002, 0.005, 5.0, 20, radii, seed=0, iteration=600)
        except RuntimeError as e:
            out[rank] = "diverged" if "diverged" in str(e) else repr(e)
            return
        sums = torch.tensor([float(p.detach().double().sum()) for p in (gm._xyz, gm._scaling, gm._features_rest)] +
                            [float(info["P"])], dtype=torch.float64)
        gathered = [torch.zeros_like(sums) for _ in range(world)]
        dist.all_gather(gathered, sums)
        out[rank] = bool(all(torch.equal(gathered[0], t) for t in gathered) and info["split"] > 0 and info["clones"] > 0)
    finally:
        dist.destroy_process_group()


def test_replicated_densification_world2():
    from tests.test_densify_emul_cpu import load_emul
    load_emul()                                            # build the emulation library before the ranks race for it
    for diverge, want in ((False, {0: True, 1: True}), (True, {0: "diverged", 1: "diverged"})):
        mgr = mp.Manager()
        out = mgr.dict()
        mp.spawn(_densify_worker, args=(2, _free_port(), out, diverge), nprocs=2, join=True)
        assert dict(out) == want


def test_camera_partition():
    for G in (1, 2, 4, 8):
        seen = sorted(sum((cameras_for_rank(8, r, G) for r in range(G)), []))
        assert seen == list(range(8))
    assert cameras_for_rank(8, 1, 4) == [1, 5]


def test_single_process_is_identity():
    p = torch.nn.Parameter(torch.ones(5, 3))
    p.grad = torch.full((5, 3), 2.0)
    fg = FlatGradients([p])
    assert fg.all_reduce() is None
    assert torch.equal(p.grad, torch.full((5, 3), 2.0))
