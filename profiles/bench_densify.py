#!/usr/bin/env python
"""Density control and Gaussian-file kernels on one B200.

  ours      gaussian_splatting_b200.densify.densify_and_prune (plan: 3 small kernels + one host wait; apply: ONE launch
            moving 6 parameter tensors + 12 Adam moments) and plyio.pack_rows / load path
  torch     the same operation in the reference's formulation (scene/gaussian_model.py:400-621): boolean-mask gathers,
            torch.cat per tensor and per moment, clone -> split -> prune as three passes over the model.  The reference's
            own class cannot travel to the GPU box (/root/reference is absent there), so this is a port of its op
            sequence; checked here against ours on the same noise before timing.

Per model size P (SH degree 3, both moments): ms per call (CUDA events around the whole call, host wait included —
it is part of the operation), the stage split of ours (plan / apply from the library's launch counter and events),
algorithmic bytes (25 P + 13 P for the plan; 708 B read per source Gaussian + 708 B written per result row for the
apply) and the fraction of the measured HBM copy peak (MEASURED_PEAKS.json, else 6552 GB/s).
Usage (GPU box): python profiles/bench_densify.py > gpurun_out/densify.json
"""
import json
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
from torch import nn  # noqa: E402

GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
ATTR = {"xyz": "_xyz", "f_dc": "_features_dc", "f_rest": "_features_rest", "opacity": "_opacity", "scaling": "_scaling",
        "rotation": "_rotation"}
SHAPES = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (3,), "rotation": (4,)}
FLOATS = 59


def hbm_peak():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        for k in ("hbm_gbs", "hbm_copy_gbps_burst", "hbm_gbps", "hbm_copy_gbps"):
            if k in d:
                return float(d[k]), f"MEASURED_PEAKS.json:{k}"
    except Exception:  # noqa: BLE001
        pass
    return 6552.0, "fallback (B200_PROFILING.md measured copy bandwidth)"


def make_state(P, dev, seed=0):
    g = torch.Generator(device=dev).manual_seed(seed)
    r = lambda *s: torch.randn(*s, device=dev, generator=g)
    s = {n: r(P, *SHAPES[n]) for n in GROUPS}
    s["opacity"] = s["opacity"] * 2.5 - 2
    s["scaling"] = s["scaling"] * 0.9 - 3
    for n in GROUPS:
        s["m_" + n], s["v_" + n] = r(P, *SHAPES[n]), torch.rand(P, *SHAPES[n], device=dev, generator=g)
    s["denom"] = torch.randint(0, 4, (P, 1), device=dev, generator=g).float()
    s["accum"] = torch.rand(P, 1, device=dev, generator=g) * 0.0006 * s["denom"]
    s["noise"] = r(2, P, 3)
    return s


def model_from(s):
    gm = types.SimpleNamespace(percent_dense=0.01)
    groups = []
    for n in GROUPS:
        p = nn.Parameter(s[n].clone())
        setattr(gm, ATTR[n], p)
        groups.append({"params": [p], "name": n})
    gm.optimizer = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
    for n in GROUPS:
        gm.optimizer.state[getattr(gm, ATTR[n])] = {"step": torch.tensor(1.0), "exp_avg": s["m_" + n].clone(),
                                                    "exp_avg_sq": s["v_" + n].clone()}
    gm.xyz_gradient_accum, gm.denom = s["accum"].clone(), s["denom"].clone()
    gm.max_radii2D = torch.zeros(s["xyz"].shape[0], device=s["xyz"].device)
    return gm


def torch_formulation(s, max_grad=0.0002, min_opacity=0.005, extent=5.0, pd=0.01):
    """clone -> split -> prune with masks, cat and index, on parameters and both moments (the reference's op sequence)."""
    t = {k: s[k] for k in s if k not in ("denom", "accum", "noise")}
    grads = torch.nan_to_num(s["accum"] / s["denom"], nan=0.0)[:, 0]
    P0 = grads.shape[0]

    def extend(new):
        for n in GROUPS:
            t[n] = torch.cat((t[n], new[n]), 0)
            t["m_" + n] = torch.cat((t["m_" + n], torch.zeros_like(new[n])), 0)
            t["v_" + n] = torch.cat((t["v_" + n], torch.zeros_like(new[n])), 0)

    def prune(mask):
        keep = ~mask
        for k in list(t):
            t[k] = t[k][keep]

    sel = (grads >= max_grad) & (torch.exp(t["scaling"]).max(1).values <= pd * extent)
    extend({n: t[n][sel] for n in GROUPS})
    padded = torch.zeros(t["xyz"].shape[0], device=grads.device)
    padded[:P0] = grads
    scal = torch.exp(t["scaling"])
    sel = (padded >= max_grad) & (scal.max(1).values > pd * extent)
    idx = sel.nonzero()[:, 0]
    stds = scal[sel].repeat(2, 1)
    samples = stds * torch.cat((s["noise"][0][idx[idx < P0]], s["noise"][1][idx[idx < P0]]), 0)
    q = torch.nn.functional.normalize(t["rotation"][sel])
    w, x, y, z = q.unbind(1)
    R = torch.stack((1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                     2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                     2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)), 1).reshape(-1, 3, 3).repeat(2, 1, 1)
    new = {n: t[n][sel].repeat(2, *([1] * (t[n].dim() - 1))) for n in GROUPS}
    new["xyz"] = torch.bmm(R, samples.unsqueeze(-1)).squeeze(-1) + new["xyz"]
    new["scaling"] = torch.log(stds / 1.6)
    extend(new)
    prune(torch.cat((sel, torch.zeros(2 * int(sel.sum()), dtype=torch.bool, device=sel.device))))
    mask = (torch.sigmoid(t["opacity"]) < min_opacity)[:, 0] | (torch.exp(t["scaling"]).max(1).values > 0.1 * extent)
    prune(mask)
    return t


def timed(fn, setup, reps):
    ms = []
    for _ in range(reps + 1):
        arg = setup()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(arg)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms = sorted(ms[1:])                                  # first call = warm-up
    return ms[len(ms) // 2]


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    from gaussian_splatting_b200 import _lib, densify, plyio
    peak, peak_src = hbm_peak()
    out = {"hbm_peak_gbps": peak, "hbm_peak_source": peak_src, "sizes": []}
    for P in (500_000, 1_000_000, 3_000_000, 6_000_000):
        s = make_state(P, dev)
        # agreement first
        gm = model_from(s)
        info = densify.densify_and_prune(gm, 0.0002, 0.005, 5.0, 20, None, noise=s["noise"])
        want = torch_formulation(s)
        same = all(gm_t.shape == want[n].shape and torch.allclose(gm_t.detach(), want[n], rtol=1e-5, atol=1e-5)
                   for n in GROUPS for gm_t in (getattr(gm, ATTR[n]),))
        l0 = _lib.launch_count()
        densify.densify_and_prune(model_from(s), 0.0002, 0.005, 5.0, 20, None, noise=s["noise"])
        launches = _lib.launch_count() - l0
        ms_ours = timed(lambda m: densify.densify_and_prune(m, 0.0002, 0.005, 5.0, 20, None, noise=s["noise"]),
                        lambda: model_from(s), 5)
        ms_torch = timed(lambda _: torch_formulation(s), lambda: None, 5)
        bytes_apply = 4 * FLOATS * 3 * (P + info["P"])
        bytes_plan = 38 * P
        gm = model_from(s)
        ms_pack = timed(lambda m: plyio.pack_rows(m), lambda: gm, 10)
        out["sizes"].append({
            "P": P, "P_new": info["P"], "kept": info["kept"], "clones": info["clones"], "split": info["split"],
            "agrees_with_torch_formulation": bool(same), "ours_ms": round(ms_ours, 4), "torch_ms": round(ms_torch, 4),
            "speedup": round(ms_torch / ms_ours, 2), "ours_kernel_launches": int(launches),
            "algorithmic_bytes": bytes_apply + bytes_plan,
            "ours_gbps_whole_call": round((bytes_apply + bytes_plan) / ms_ours / 1e6, 1),
            "frac_of_hbm_peak_whole_call": round((bytes_apply + bytes_plan) / ms_ours / 1e6 / peak, 3),
            "ply_pack_ms": round(ms_pack, 4), "ply_pack_gbps": round(2 * 4 * 62 * P / ms_pack / 1e6, 1)})
        del s, gm, want
        torch.cuda.empty_cache()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
