#!/usr/bin/env python
"""Benchmark of the hot path: fwd+bwd Gaussian rasterization (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C3]

One "step" = one forward + backward pass of the rasterizer over one camera of the synthetic
workload (default C3 = 3M Gaussians, 1920x1080, SH degree 3, antialiasing on,
dense upstream gradients for colour and inverse depth).  With N > 1 (torchrun, one rank per
GPU) every rank rasterizes its own camera of a ring and the per-Gaussian parameter gradients
are summed with one NCCL all-reduce per step (view-parallel, weak scaling).

Prints ONE JSON line on rank 0.  Keys beyond the base contract:
  roofline      dominant stage of the step, timed live with CUDA events on the launching stream
  stages        every stage: ms, algorithmic bytes / flops, achieved, peak, fraction
  work          measured work descriptors P, V, R, T, N, Wf, Wb, Bp (independent counting pass)
  cpu_baseline  the CPU oracle (port of the reference kernels) on the host cores
  e2e           same metric through the public API with host-side step inputs (camera + targets
                copied from pinned memory every step, loss read back every step)
  fwd_fps       forward-only render FPS (second half of BASELINE.json's metric)
`--impl reference` times the reference's own CUDA rasterizer (oracle/_ref, built from
/root/reference by oracle/build_ref.py) with the same protocol; if that build is absent it
falls back to the CPU oracle port and says so.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

METRIC = "fwd+bwd rasterize iters/sec at 3M Gaussians 1080p"
UNIT = "iter/s"
SM_COUNT_B200 = 148
FP32_LANES_PER_SM = 128


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)      # 10 warm-up, >= 50 timed iterations
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C3", help="C1|C2|C3|C4 (SURVEY.md §8d) or C3 with --P override")
    ap.add_argument("--P", type=int, default=None, help="override the Gaussian count")
    ap.add_argument("--yaw-step", type=float, default=45.0,
                    help="rank r looks at the scene from yaw r * step degrees (0: every rank has the same camera; diagnostic)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--ref-variant", default="restored", choices=["restored", "asis"])
    ap.add_argument("--reduce", default="fused", choices=["fused", "nccl"],
                    help="N > 1: gradient reduction inside the backward kernel over NVLink multicast (fused) "
                         "or a separate NCCL all-reduce of the packed gradients (nccl)")
    return ap.parse_args()


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm_gbs=float(d["hbm_gbs"]), sm_max_mhz=float(d.get("sm_max_mhz", 1965.0)), source="measured")
    return dict(hbm_gbs=6650.0, sm_max_mhz=1965.0, source="fallback")


# ------------------------------------------------------------------ clocks during the timed region
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        for ts, line in self.rows:
            if t0 is not None and not (t0 - 0.05 <= ts <= t1 + 0.15):
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["no samples"])
        sm.sort()
        return dict(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(smax), reasons=sorted(reasons),
                    samples=len(sm), power_w_max=max(power))


# ------------------------------------------------------------------ workload
def build_workload(args, rank, device):
    """Synthetic Gaussians (identical on every rank) + this rank's camera + upstream gradients."""
    from gaussian_splatting_b200 import synthetic
    cfg = synthetic.CONFIGS[args.workload]
    P = args.P or cfg.P
    gs, cam = synthetic.make_scene(cfg, P=P, yaw_deg=args.yaw_step * rank)
    W, H = cfg.W, cfg.H
    g = torch.Generator().manual_seed(1000 + cfg.seed)
    target, target_d = synthetic.make_targets(W, H, cfg.seed)
    dc = (torch.rand(3, H, W, generator=g) - 0.5) / (3 * H * W)
    dd = (torch.rand(1, H, W, generator=g) - 0.5) * 0.1 / (H * W)
    s = dict(cam)
    s.update(bg=torch.zeros(3), scale_modifier=1.0, sh_degree=3, prefiltered=False, debug=False)
    wl = dict(cfg=cfg, P=P, W=W, H=H, gs=gs, s=s, dc=dc, dd=dd, target=target, target_d=target_d)
    if device is not None:
        wl["params"] = {k: v.to(device).requires_grad_(True) for k, v in gs.items()}
        wl["means2D"] = torch.zeros(P, 3, device=device, requires_grad=True)
        wl["dc_dev"], wl["dd_dev"] = dc.to(device), dd.to(device)
    return wl


def settings_of(module, s, device, cam_dev=None):
    cam_dev = cam_dev or {}

    def dev(k):          # NOT dict.get(k, s[k].to(device)): the default would be evaluated every time, and a pageable
        return cam_dev[k] if k in cam_dev else s[k].to(device)   # host-to-device copy synchronises the stream

    return module.GaussianRasterizationSettings(
        image_height=int(s["image_height"]), image_width=int(s["image_width"]),
        tanfovx=float(s["tanfovx"]), tanfovy=float(s["tanfovy"]), bg=dev("bg"),
        scale_modifier=1.0, viewmatrix=dev("viewmatrix"), projmatrix=dev("projmatrix"), sh_degree=3,
        campos=dev("campos"), prefiltered=False, debug=False, antialiasing=bool(s["antialiasing"]))


def make_steps(module, wl, device, flat=None):
    """Returns (step_resident, step_forward_only, step_e2e, e2e_bytes)."""
    params, means2D = wl["params"], wl["means2D"]
    leaves = list(params.values()) + [means2D]
    rs = settings_of(module, wl["s"], device)
    rast = module.GaussianRasterizer(rs)
    dc, dd = wl["dc_dev"], wl["dd_dev"]

    def clear():
        for t in leaves:
            t.grad = None

    def step_resident():
        clear()
        color, radii, invd = rast(means3D=params["means3D"], means2D=means2D, opacities=params["opacities"],
                                  shs=params["shs"], scales=params["scales"], rotations=params["rotations"])
        torch.autograd.backward([color, invd], [dc, dd])
        if flat is not None:
            flat.all_reduce()

    def step_forward_only():
        with torch.no_grad():
            rast(means3D=params["means3D"], means2D=means2D, opacities=params["opacities"],
                 shs=params["shs"], scales=params["scales"], rotations=params["rotations"])

    # ---- end to end: this step's camera and targets come from pinned host memory, loss goes back
    H, W = wl["H"], wl["W"]
    cam_host = torch.cat([wl["s"]["viewmatrix"].reshape(-1), wl["s"]["projmatrix"].reshape(-1),
                          wl["s"]["campos"].reshape(-1)]).pin_memory()
    tgt_host = torch.cat([wl["target"].reshape(-1), wl["target_d"].reshape(-1)]).pin_memory()
    cam_dev = torch.empty_like(cam_host, device=device)
    tgt_dev = [torch.empty_like(tgt_host, device=device) for _ in range(2)]   # double-buffered: step i+1's targets
    copied = [torch.cuda.Event(), torch.cuda.Event()]                         # land while step i is computed
    consumed = [torch.cuda.Event(), torch.cuda.Event()]
    loss_host = torch.zeros(2).pin_memory()      # ring of two slots: step i's loss is read while step i+1 is enqueued
    loss_ready = [torch.cuda.Event(), torch.cuda.Event()]
    pending = []                                 # [(slot)] of the step whose loss has not been read yet
    losses = []
    bg_dev = wl["s"]["bg"].to(device)
    h2d = cam_host.numel() * 4 + tgt_host.numel() * 4
    d2h = 4

    copy_stream = torch.cuda.Stream(device=device)
    counter = [0]

    def fetch_targets(buf):
        """One pinned-host -> device copy of a step's target image / depth on the copy stream."""
        copy_stream.wait_event(consumed[buf])                # the step that last used this buffer has read it
        with torch.cuda.stream(copy_stream):
            tgt_dev[buf].copy_(tgt_host, non_blocking=True)
            copied[buf].record(copy_stream)

    def drain():
        while pending:
            slot = pending.pop(0)
            loss_ready[slot].synchronize()
            losses.append(float(loss_host[slot]))

    def l1_losses(color, invd, tgt, tgt_d):
        return (color - tgt).abs().mean() + 0.1 * (invd - tgt_d).abs().mean()

    def step_e2e(loss_fn=l1_losses):
        clear()
        main = torch.cuda.current_stream(device)
        cam_dev.copy_(cam_host, non_blocking=True)          # 140 B: needed before the first kernel
        # the 33 MB of target image / depth are only needed by the loss.  Like a prefetching data loader, every step
        # issues the copy of the NEXT step's targets (other buffer, copy stream), so each step makes exactly one such
        # copy inside the timed region and a slow PCIe link does not stall the forward-to-loss hand-over.
        buf = counter[0] & 1
        if counter[0] == 0:
            fetch_targets(buf)                               # the very first (warm-up) step fetches its own
        fetch_targets(buf ^ 1)
        cd = dict(viewmatrix=cam_dev[:16].view(4, 4), projmatrix=cam_dev[16:32].view(4, 4), campos=cam_dev[32:35],
                  bg=bg_dev)
        r = module.GaussianRasterizer(settings_of(module, wl["s"], device, cd))
        color, radii, invd = r(means3D=params["means3D"], means2D=means2D, opacities=params["opacities"],
                               shs=params["shs"], scales=params["scales"], rotations=params["rotations"])
        main.wait_event(copied[buf])
        tgt = tgt_dev[buf][:3 * H * W].view(3, H, W)
        tgt_d = tgt_dev[buf][3 * H * W:].view(1, H, W)
        loss = loss_fn(color, invd, tgt, tgt_d)
        loss.backward()
        consumed[buf].record(main)
        if flat is not None:
            flat.all_reduce()
        # Every step's loss crosses to the host; it is READ one step late (event-guarded pinned slot), so the host
        # enqueues step i+1 while the device finishes step i instead of idling the device during the host's set-up.
        slot = counter[0] & 1
        counter[0] += 1
        loss_host[slot:slot + 1].copy_(loss.detach().reshape(1), non_blocking=True)
        loss_ready[slot].record(torch.cuda.current_stream(device))
        pending.append(slot)
        if len(pending) > 1:
            s0 = pending.pop(0)
            loss_ready[s0].synchronize()
            losses.append(float(loss_host[s0]))

    step_e2e.drain = drain
    step_e2e.losses = losses
    return step_resident, step_forward_only, step_e2e, (h2d, d2h)


def timed_loop(step, steps, warmup, world, device, on_start=None, per_step=None, drain=None):
    """W untimed steps, then exactly K steps between barrier + synchronize; device time by CUDA
    events on the launching stream, max over ranks.  Returns (ms_total_max, wall t0, wall t1).
    per_step: list that receives this rank's K step durations (an event after every step).
    drain: called after the last step, inside the timed region (a pipelined step's pending read-back)."""
    for _ in range(warmup):
        step()
    if drain is not None:
        drain()
    torch.cuda.synchronize(device)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    if on_start is not None:
        on_start()
    t0 = time.time()
    ev[0].record()
    for i in range(steps):
        step()
        ev[i + 1].record()
    if drain is not None:
        drain()
    torch.cuda.synchronize(device)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    t1 = time.time()
    ms = ev[0].elapsed_time(ev[-1])
    if per_step is not None:
        per_step.extend(ev[i].elapsed_time(ev[i + 1]) for i in range(steps))
    if world > 1:
        t = torch.tensor([ms], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
    return ms, t0, t1


def percentiles(ms_list):
    """median / p10 / p90 of the per-step device times."""
    if not ms_list:
        return None
    v = sorted(ms_list)
    q = lambda f: v[min(len(v) - 1, max(0, int(round(f * (len(v) - 1)))))]  # noqa: E731
    return {"median": round(q(0.5), 4), "p10": round(q(0.1), 4), "p90": round(q(0.9), 4), "n": len(v)}


def measure_fp32_peak(device):
    """Scalar-FFMA rate of this GPU at its clocks right now (gsr_debug_fp32_peak), TFLOP/s; best of 5 launches."""
    import ctypes as C
    from gaussian_splatting_b200 import _lib
    lib = _lib.load()
    sink = torch.zeros(1, device=device)
    flops = C.c_double(0.0)
    st = torch.cuda.current_stream(device).cuda_stream
    best = 0.0
    for i in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(lib.gsr_debug_fp32_peak(SM_COUNT_B200 * 16, 4096, sink.data_ptr(), C.byref(flops), st))
        e1.record()
        e1.synchronize()
        if i:
            best = max(best, flops.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best


# ------------------------------------------------------------------ roofline bookkeeping
def stage_model(work):
    """Algorithmic bytes / flops of each stage of OUR path for one step.
    HBM stages: minimum bytes the stage must move; render stages: FLOP model of SURVEY.md §8d."""
    P, V, R, T, N = work["P"], work["V"], work["R"], work["T"], work["N"]
    Wf, Wb, Bp = work["Wf"], work["Wb"], work["Bp"]
    C = work.get("chunks", 1)            # chunks of depth ranks in the tile-list construction (common.cuh: tile_bin_plan)
    world = work.get("world", 1)
    m = {
        # means 12 + radii 4 + depth key 8 + rect 8 for all; scales/rot/opacity/SH 224 read + record 64 written for visible
        "preprocess": ("hbm", 32 * P + 288 * V),
        # one histogram read (8 P) + 4 LSD passes that read and write the 8-byte items once each (16 P)
        "depth_sort": ("hbm", 72 * P),
        # key 8 + rect gather 8 + depth-ordered rect 8 per visible Gaussian; the (chunk, tile) counts written (4 C T),
        # read and rewritten as prefixes (8 C T); totals 4 T written + read, ranges 8 T written
        "tile_count": ("hbm", 24 * V + 12 * C * T + 16 * T),
        # key 8 + rect 8 per visible Gaussian, the chunk prefixes 4 C T + tile starts 8 T read, 4 bytes per instance written
        "tile_fill": ("hbm", 16 * V + 4 * C * T + 8 * T + 4 * R),
        "render_fwd": ("fp32", 14 * Wf + 11 * Bp),
        "render_bwd": ("fp32", 14 * Wb + 75 * Bp),
        # dense outputs for every Gaussian + visible reads (+ 48 B accumulator)
        "preprocess_bwd": ("hbm", 236 * P + 287 * V + 48 * V),
    }
    if world > 1:
        # factored exchange: the sender branch writes dL_dmeans2D densely (12 P) and publishes 64 B per contributing
        # Gaussian instead of 236 B rows; the expand pass reads one 16 B colour-gradient slot per rank, the 48 B
        # record and the mean of every Gaussian and writes the 236 B dense gradient row
        m["preprocess_bwd"] = ("hbm", 12 * P + (287 + 48 + 64) * V)
        m["exchange"] = ("hbm", (16 * world + 48 + 12 + 236) * P)
    return m


def rooflines(stage_ms, work, peaks, sm_mhz):
    model = stage_model(work)
    # measured scalar-FFMA rate when available (gsr_debug_fp32_peak), else the nominal CUDA-core peak at max clock
    fp32_peak = peaks.get("fp32_tflops") or SM_COUNT_B200 * FP32_LANES_PER_SM * 2 * peaks["sm_max_mhz"] * 1e6 / 1e12
    out = []
    for name, (ms_total, calls) in stage_ms.items():
        if not calls or name not in model:
            continue
        ms = ms_total / calls
        bound, amount = model[name]
        if bound == "hbm":
            achieved = amount / (ms * 1e-3) / 1e9
            peak, unit = peaks["hbm_gbs"], "GB/s"
        else:
            achieved = amount / (ms * 1e-3) / 1e12
            peak, unit = fp32_peak, "TFLOP/s"
        out.append(dict(stage=name, ms=round(ms, 4), bound=bound, algorithmic=int(amount), achieved=round(achieved, 3),
                        peak=round(peak, 1), unit=unit, frac=round(achieved / peak, 4)))
    out.sort(key=lambda d: -d["ms"])
    return out


STAGE_KERNELS = {"render_bwd": ("render_bwd_kernel", "render_bwd.cu"), "render_fwd": ("render_fwd_kernel", "render_fwd.cu"),
                 "preprocess_bwd": ("preprocess_bwd_kernel", "preprocess_bwd.cu"),
                 "preprocess": ("preprocess_fwd_kernel", "preprocess.cu")}


def source_sha1(src):
    import hashlib
    return hashlib.sha1(open(os.path.join(ROOT, "gaussian_splatting_b200", "csrc", src), "rb").read()).hexdigest()[:16]


def ncu_evidence(stage):
    """DRAM traffic per launch and issue-slot utilisation of a stage's kernel from a committed `ncu --set full`
    capture of this same workload (profiles/**/r*_kernels*.json; profiles/ncu_kernels_json.py stamps every entry with
    the sha1 of the kernel's source file at capture time).  Returns None when there is no capture of the kernel as it
    is now, so that a stale number never rides on a fresh line."""
    import glob
    if stage not in STAGE_KERNELS:
        return None
    kernel, src = STAGE_KERNELS[stage]
    now = source_sha1(src)
    import re
    # rNN<letter>_kernels[_fwd|_bwd].json are captures of the bench workload (profiles/gpu_round.sh); the *_misc_* and
    # *_densify_* files hold other workloads' launches of the same kernels and must not be matched
    found = sorted(f for f in glob.glob(os.path.join(ROOT, "profiles", "**", "r*_kernels*.json"), recursive=True)
                   if re.fullmatch(r"r\d+[a-z]*_kernels(_fwd|_bwd)?\.json", os.path.basename(f)))
    for path in reversed(found):
        try:
            entries = json.load(open(path))
        except Exception:   # noqa: BLE001
            continue
        for k in entries:
            if k.get("kernel", "").startswith(kernel) and k.get("src_sha1") == now:
                return dict(traffic=int((k["dram_read_MB"] + k["dram_write_MB"]) * 1e6),
                            issue_active=round(k["issue_active_pct"] / 100.0, 3),
                            source="profiles/" + os.path.relpath(path, os.path.join(ROOT, "profiles")),
                            src_sha1=now)
    return None


def measure_work(wl, device):
    """Work descriptors of this rank's view from the state of one forward (independent counting pass)."""
    from gaussian_splatting_b200 import rasterizer
    import diff_gaussian_rasterization as dgr
    rs = settings_of(dgr, wl["s"], device)
    p = wl["params"]
    e = torch.Tensor([])
    st = {}
    with torch.no_grad():
        R, color, radii, *_ = rasterizer._forward_native(rs, p["means3D"], p["shs"], e, p["opacities"], p["scales"],
                                                         p["rotations"], e, state_out=st)
    c = st["work_counts"].cpu().tolist()
    W, H = wl["W"], wl["H"]
    gx, gy = (W + 15) // 16, (H + 15) // 16
    band_rows = max(1, min(gy, 512, 8192 // gx))                       # common.cuh: tile_bin_plan
    chunks = max(1, min(888, (32 << 20) // (gx * gy), wl["P"] // 64))
    return dict(P=wl["P"], V=int((radii > 0).sum()), R=int(R), T=gx * gy, N=W * H,
                Wf=int(c[0]), Bp=int(c[1]), Wb=int(c[2]), chunks=chunks)


# ------------------------------------------------------------------ CPU baseline (oracle port)
def cpu_oracle_iteration(wl, threads, reps=1):
    """One full fwd+bwd of the CPU oracle (C restatement of the reference kernels) on the host.
    Returns (seconds per iteration, description)."""
    from oracle import cpu_ref
    cpu_ref.build()
    cpu_ref.set_threads(threads)
    gs = {k: v.numpy() for k, v in wl["gs"].items()}
    s = {k: (v.numpy() if isinstance(v, torch.Tensor) else v) for k, v in wl["s"].items()}
    dc, dd = wl["dc"].numpy(), wl["dd"].numpy().reshape(-1)
    best = None
    for _ in range(reps):
        t0 = time.time()
        fw = cpu_ref.forward(gs, s)
        cpu_ref.backward(fw, s, dc, dd, restored=True)
        dt = time.time() - t0
        best = dt if best is None else min(best, dt)
    return best, f"full workload ({wl['cfg'].name}: P={wl['P']}, {wl['W']}x{wl['H']}), fwd+bwd, best of {reps}"


# ------------------------------------------------------------------ arms
def init_dist(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def base_line(args, wl, world, impl):
    return {
        "metric": METRIC, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "impl": impl,
        "config": {"workload": f"{wl['cfg'].name}: {wl['P']} Gaussians, {wl['W']}x{wl['H']}, SH degree 3, "
                               f"antialiasing={bool(wl['s']['antialiasing'])}, fwd+bwd with colour and "
                               "inverse-depth gradients, one camera per GPU per step",
                   "parallelism": f"view-parallel x{world}" if world > 1 else "single GPU",
                   "l2_policy": f"inputs larger than L2 ({wl['P'] * 236 / 1e6:.0f} MB of Gaussian parameters read per step "
                                "vs 126 MB L2); no explicit flush",
                   "seed": wl["cfg"].seed},
    }


def reduce_check(module, wl, device, world, flat=None):
    """N > 1: the gradients the step's reduction delivers (the fused exchange inside the backward kernel, or the packed
    all-reduce of FlatGradients with --reduce nccl) against the same gradients computed per rank without it and summed
    tensor by tensor with dist.all_reduce.  Raises above 1e-5 (norm-wise), so a SCALE line only exists for correct sums."""
    from gaussian_splatting_b200 import rasterizer as rz
    params, means2D = wl["params"], wl["means2D"]
    rast = module.GaussianRasterizer(settings_of(module, wl["s"], device))
    dc, dd = wl["dc_dev"], wl["dd_dev"]

    def grads():
        nonlocal use_flat
        for t in list(params.values()) + [means2D]:
            t.grad = None
        color, radii, invd = rast(means3D=params["means3D"], means2D=means2D, opacities=params["opacities"],
                                  shs=params["shs"], scales=params["scales"], rotations=params["rotations"])
        torch.autograd.backward([color, invd], [dc, dd])
        if use_flat and flat is not None:
            flat.all_reduce()
        torch.cuda.synchronize(device)
        return {k: v.grad.detach().clone() for k, v in params.items()}

    use_flat = True
    fused = grads()
    use_flat = False
    red = rz._gradient_reducer
    rz.set_gradient_reducer(None)
    try:
        plain = grads()
    finally:
        rz.set_gradient_reducer(red)
    worst = 0.0
    per = {}
    for k, g in plain.items():
        dist.all_reduce(g, op=dist.ReduceOp.SUM)
        err = float((fused[k] - g).abs().max() / g.abs().max().clamp_min(1e-30))
        per[k] = float(f"{err:.3e}")
        worst = max(worst, err)
    t = torch.tensor([worst], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    worst = float(t[0])
    if not worst < 1e-5:
        raise SystemExit(f"reduce_check failed: fused exchange differs from all_reduce by {worst:.3e} ({per})")
    return {"max_rel_err": float(f"{worst:.3e}"), "ranks": world, "per_tensor_rank0": per,
            "what": "||fused - all_reduce(unfused)||inf / ||all_reduce(unfused)||inf, max over tensors and ranks"}


def traced_gaps(step, device, n=10):
    """Mean device-time gaps between the phase boundaries of a step on this rank (events on the launching stream):
    bwd0->begun = wait for the all-ranks-cleared barrier, begun->sent = blend backward + publishing kernel,
    sent->barrier = all-ranks-published barrier, barrier->expanded = gsr_exchange_expand."""
    from gaussian_splatting_b200 import rasterizer as rz
    was = rz._TRACE
    rz._TRACE = True
    try:
        rz.trace_reset()
        for _ in range(n):
            step()
        torch.cuda.synchronize(device)
        return rz.trace_summary()
    finally:
        rz._TRACE = was
        rz.trace_reset()


def run_gpu_arm(args, module, impl, rank, world, local):
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=device)
    wl = build_workload(args, rank, device)
    flat = None
    reduce_mode = None
    if world > 1:
        from gaussian_splatting_b200.view_parallel import FlatGradients, FusedGradientReducer
        reduce_mode = "nccl all-reduce of the packed gradients"
        if impl == "ours" and args.reduce == "fused" and FusedGradientReducer.available():
            try:
                red = FusedGradientReducer().install()
                reduce_mode = ("fused: factored exchange (64 B per contributing Gaussian via multimem.red into NVLink-multicast "
                               "buffers inside the backward kernel, SH gradients rebuilt locally per sender camera)"
                               if red.factored else
                               f"fused: multimem.red of all gradients into NVLink-multicast buffers inside the backward kernel, "
                               f"{red.bands} band(s)")
            except Exception as ex:      # noqa: BLE001
                reduce_mode += f" (fused unavailable: {ex})"
        if not reduce_mode.startswith("fused"):
            flat = FlatGradients(list(wl["params"].values()))
    step, step_fwd, step_e2e, (h2d, d2h) = make_steps(module, wl, device, flat)
    ours = impl == "ours"
    if ours:
        from gaussian_splatting_b200 import _lib
        work = measure_work(wl, device)
        work["world"] = world

    sampler = ClockSampler(local)
    sampler.start()
    mark = {}

    def on_start():
        if ours:
            _lib.profile_enable(True)
            mark["l0"] = _lib.launch_count()

    from gaussian_splatting_b200 import rasterizer as _rz
    def on_start_all():
        on_start()
        _rz.trace_reset()
    per_step = []
    ms, t0, t1 = timed_loop(step, args.steps, args.warmup, world, device, on_start_all, per_step=per_step)
    trace = _rz.trace_summary() if ours else {}
    clocks = sampler.stop(t0, t1)
    line = base_line(args, wl, world, impl)
    if reduce_mode:
        line["config"]["gradient_reduction"] = reduce_mode
    if ours:
        launches = _lib.launch_count() - mark["l0"]
        stage_ms = _lib.profile_collect()
        _lib.profile_enable(False)
    ms_step = ms / args.steps
    line.update(value=round(world * 1000.0 / ms_step, 3), ms_per_step=round(ms_step, 4), clocks=clocks)
    line["ms_per_step_percentiles"] = percentiles(per_step)      # this rank's steps (median / p10 / p90)

    # forward-only FPS
    ms_f, _, _ = timed_loop(step_fwd, args.steps, 3, world, device)
    line["fwd_fps"] = round(world * 1000.0 * args.steps / ms_f, 2)

    if not args.no_e2e:
        per_e = []
        ms_e, _, _ = timed_loop(step_e2e, args.steps, 3, world, device, per_step=per_e, drain=step_e2e.drain)
        line["e2e"] = {"value": round(world * 1000.0 * args.steps / ms_e, 3), "unit": UNIT,
                       "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                       "ms_per_step_percentiles": percentiles(per_e),
                       "what": "camera + target image/depth from pinned host memory (targets double-buffered on a copy stream: every step issues the copy of the next step's targets, one 33 MB copy per step inside the timed region), L1 losses in torch, "
                               "backward, loss scalar copied to pinned memory every step and read by the host one step "
                               "late (event-guarded ring of two slots; the last one is drained inside the timed region); "
                               "Gaussian parameters stay resident (they are the model)"}
    if ours and not args.no_e2e and world == 1:
        # second figure (this repo only; the arms' common protocol above stays the headline): the same end-to-end step
        # with the FULL loss of train.py:155-174 (0.8 L1 + 0.2 (1 - SSIM) + 0.1 inverse-depth L1), once as the
        # reference's train.py would compute it on top of this repo's drop-in packages (torch L1 terms + fused_ssim)
        # and once as one fused forward + one fused backward launch (gaussian_splatting_b200.loss)
        import fused_ssim as _fs
        from gaussian_splatting_b200 import loss as _loss

        def full_unfused(color, invd, tgt, tgt_d):
            return 0.8 * (color - tgt).abs().mean() + 0.2 * (1.0 - _fs.fused_ssim(color.unsqueeze(0), tgt.unsqueeze(0))) \
                + 0.1 * (invd - tgt_d).abs().mean()

        def full_fused(color, invd, tgt, tgt_d):
            return _loss.training_loss(color, tgt, 0.2, invdepth=invd, mono_invdepth=tgt_d, depth_weight=0.1)

        fig = {}
        for name, fn in (("torch_l1_plus_fused_ssim", full_unfused), ("fused_training_loss", full_fused)):
            ms_x, _, _ = timed_loop(lambda: step_e2e(fn), args.steps, 3, world, device, drain=step_e2e.drain)
            fig[name] = round(1000.0 * args.steps / ms_x, 3)
        fig["unit"] = UNIT
        fig["what"] = "e2e protocol with the full loss of train.py:155-174 (lambda_dssim 0.2, depth weight 0.1)"
        line["e2e_full_loss"] = fig
    if ours and world > 1:
        line["reduce_check"] = reduce_check(module, wl, device, world, flat)
        line["trace_ms"] = traced_gaps(step, device)
    if ours:
        peaks = load_peaks()
        peaks["fp32_tflops"] = round(measure_fp32_peak(device), 2)
        stages = rooflines(stage_ms, work, peaks, clocks.get("sm_mhz"))
        line["gpu_launches"] = int(launches)
        if trace and "trace_ms" not in line:
            line["trace_ms"] = trace          # GSR_TRACE_EVENTS=1: mean gaps between phase boundaries on this rank
        line["work"] = work
        line["stages"] = stages
        if world > 1:
            # every rank's own kernel time per step (sum of its stages): separates the work imbalance between
            # the cameras of the ring from waiting at the two rank barriers (value uses the max over ranks)
            mine = torch.tensor([sum(s["ms"] for s in stages), float(work["R"])], device=device, dtype=torch.float64)
            allr = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(allr, mine)
            line["per_rank"] = {"kernel_ms_per_step": [round(float(t[0]), 3) for t in allr],
                                "num_rendered": [int(t[1]) for t in allr]}
        top = stages[0]
        line["roofline"] = {"kernel": top["stage"], "bound": top["bound"], "achieved": top["achieved"],
                            "peak": top["peak"], "unit": top["unit"], "frac": top["frac"], "traffic": None,
                            "ms": top["ms"],
                            "peak_source": (f"{peaks['source']} copy bandwidth (MEASURED_PEAKS.json)" if top["bound"] == "hbm"
                                            else "scalar FFMA rate measured in this run (gsr_debug_fp32_peak, best of 5 "
                                                 "launches; nominal 148 SMs x 128 lanes x 2 x max SM clock = "
                                                 f"{SM_COUNT_B200 * FP32_LANES_PER_SM * 2 * peaks['sm_max_mhz'] / 1e6:.1f})")}
        ev = ncu_evidence(top["stage"]) if wl["cfg"].name == "C3" and wl["P"] == wl["cfg"].P else None
        if ev:
            line["roofline"].update(traffic=ev["traffic"], issue_slot_utilisation=ev["issue_active"],
                                    ncu_source=ev["source"], ncu_src_sha1=ev["src_sha1"])
        hbm = [s for s in stages if s["bound"] == "hbm"]
        if hbm:
            line["roofline_hbm"] = hbm[0]
        xch = [s for s in stages if s["stage"] == "exchange"]
        if xch:
            # N > 1: the receiver half of the gradient exchange (gsr_exchange_expand) against the HBM peak; the two
            # rank barriers and the publishing kernels are in trace_ms
            line["roofline_exchange"] = xch[0]
    else:
        line["gpu_launches"] = 0          # none of this repo's kernels run in the reference arm
    return line, wl


def run_cpu_port_arm(args, why):
    """Fallback of --impl reference when the reference CUDA build is absent: CPU oracle port."""
    wl = build_workload(args, 0, None)
    threads = min(os.cpu_count() or 1, 256)
    for _ in range(min(args.warmup, 1)):
        cpu_oracle_iteration(wl, threads)
    t = [cpu_oracle_iteration(wl, threads)[0] for _ in range(min(args.steps, 3))]
    sec = sum(t) / len(t)
    line = base_line(args, wl, 1, "reference")
    line.update(value=round(1.0 / sec, 4), ms_per_step=round(sec * 1e3, 2), steps=len(t), gpu_launches=0,
                cpu_baseline={"value": round(1.0 / sec, 4), "unit": UNIT, "cores": threads, "kind": "port",
                              "sample": f"full workload, mean of {len(t)} iterations", "why": why},
                e2e={"value": round(1.0 / sec, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
    return line


def main():
    args = parse_args()
    rank, world, local = init_dist(args)
    if args.impl == "reference":
        if rank != 0:
            return 0                     # the reference is single-GPU: rank 0 alone runs it
        from oracle import ref_loader
        why = None
        if not torch.cuda.is_available():
            why = "no CUDA device"
        elif not ref_loader.available(args.ref_variant):
            why = f"oracle/_ref/{args.ref_variant} not built"
        if why is None:
            try:
                ref = ref_loader.load(args.ref_variant)
            except Exception as ex:      # noqa: BLE001
                why = f"reference build failed to load: {ex}"
        if why is not None:
            print(json.dumps(run_cpu_port_arm(args, why)))
            return 0
        line, wl = run_gpu_arm(args, ref, "reference", 0, 1, local)
        line["n_gpus"] = 1
        # the reference ships no CPU rasterizer: its own implementation of the path is
        # CUDA, timed here on one B200; the CPU port of the same math is `cpu_baseline` of the ours arm
        line["cpu_baseline"] = {"value": line["value"], "unit": UNIT, "cores": 0, "kind": "reference",
                                "sample": "full workload on one B200 (the reference's only implementation is CUDA)"}
        line["reference"] = {"kind": "reference", "what": "the reference's own CUDA rasterizer "
                             "(submodules/diff-gaussian-rasterization) compiled for sm_100 by oracle/build_ref.py; "
                             f"variant '{args.ref_variant}' (SURVEY.md F1-F3 repairs)"}
        print(json.dumps(line))
        return 0

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the product has no CPU path)")
    import diff_gaussian_rasterization as dgr
    line, wl = run_gpu_arm(args, dgr, "ours", rank, world, local)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = min(os.cpu_count() or 1, 256)
        sec, sample = cpu_oracle_iteration(wl, threads)
        line["cpu_baseline"] = {"value": round(1.0 / sec, 4), "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": sample}
    if rank == 0:
        print(json.dumps(line))
    if world > 1 and dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
