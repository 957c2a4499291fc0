#!/usr/bin/env python
"""Device timeline of bench.py's end-to-end step (C3, one GPU) from torch.profiler (CUPTI): per-kernel and per-copy
device time per step, and the idle time of the GPU inside a step.  Answers "where do the 0.4 ms between `value` and
`e2e` go".  Usage (GPU box): python profiles/e2e_trace.py [--resident] [out.json]"""
import json
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
import diff_gaussian_rasterization as dgr  # noqa: E402

STEPS = 10
args = types.SimpleNamespace(workload="C3", P=None, yaw_step=45.0)
device = torch.device("cuda", 0)
torch.cuda.set_device(device)
wl = bench.build_workload(args, 0, device)
step_resident, _, step_e2e, _ = bench.make_steps(dgr, wl, device)
if "--resident" in sys.argv:           # the `value` loop instead (everything resident, upstream gradients given)
    sys.argv.remove("--resident")
    step_resident.drain = lambda: None
    step_e2e = step_resident
for _ in range(10):
    step_e2e()
step_e2e.drain()
torch.cuda.synchronize()
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA, torch.profiler.ProfilerActivity.CPU]) as prof:
    for _ in range(STEPS):
        step_e2e()
    step_e2e.drain()
    torch.cuda.synchronize()

ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
per = {}
for e in ev:
    k = e.name[:70]
    d = per.setdefault(k, [0, 0.0])
    d[0] += 1
    d[1] += e.device_time if hasattr(e, "device_time") else e.cuda_time
rows = sorted(per.items(), key=lambda kv: -kv[1][1])
total = 0.0
out = []
for k, (n, us) in rows:
    out.append(dict(name=k, launches_per_step=n / STEPS, us_per_step=round(us / STEPS, 1)))
    total += us / STEPS
# busy time of the device = union of the intervals of all device activities
iv = sorted((e.time_range.start, e.time_range.end) for e in ev)
busy, cur_s, cur_e = 0.0, None, None
for s, e in iv:
    if cur_e is None or s > cur_e:
        if cur_e is not None:
            busy += cur_e - cur_s
        cur_s, cur_e = s, e
    else:
        cur_e = max(cur_e, e)
if cur_e is not None:
    busy += cur_e - cur_s
span = iv[-1][1] - iv[0][0] if iv else 0.0
# idle gaps: which activity ends before the gap and which one starts after it
named = sorted(((e.time_range.start, e.time_range.end, e.name[:48]) for e in ev))
gaps = {}
last_end, last_name = None, None
for s_, e_, n_ in named:
    if last_end is not None and s_ > last_end + 1.0:
        g = gaps.setdefault((last_name, n_), [0, 0.0])
        g[0] += 1
        g[1] += s_ - last_end
    if last_end is None or e_ > last_end:
        last_end, last_name = e_, n_
gap_rows = sorted(gaps.items(), key=lambda kv: -kv[1][1])[:12]
res = dict(steps=STEPS, span_us_per_step=round(span / STEPS, 1), busy_us_per_step=round(busy / STEPS, 1),
           idle_us_per_step=round((span - busy) / STEPS, 1), sum_of_activities_us_per_step=round(total, 1),
           activities=out[:40],
           gaps=[dict(after=k[0], before=k[1], per_step=round(v[0] / STEPS, 2), us_per_step=round(v[1] / STEPS, 1))
                 for k, v in gap_rows])
path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "e2e_trace.json")
json.dump(res, open(path, "w"), indent=1)
print(json.dumps({k: v for k, v in res.items() if k not in ("activities", "gaps")}))
for g in res["gaps"]:
    print(f"gap {g['us_per_step']:7.1f} us/step x{g['per_step']:<5}  after {g['after']}  before {g['before']}")
for r in out[:32]:
    print(f"{r['us_per_step']:9.1f} us  x{r['launches_per_step']:<5.1f} {r['name']}")
