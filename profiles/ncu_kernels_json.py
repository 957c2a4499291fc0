#!/usr/bin/env python
"""`ncu -i X.ncu-rep --page raw --csv` -> compact per-launch JSON (what bench.py's ncu_evidence() reads).
Usage: python profiles/ncu_kernels_json.py raw.csv > profiles/r01g_kernels.json"""
import csv
import hashlib
import json
import os
import re
import sys

CSRC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gaussian_splatting_b200", "csrc")
# kernel-name prefix -> source file: every entry carries the sha1 of that file as it is when this script runs (run it
# right after the capture), so bench.py only quotes a capture of the kernel as it is now
SOURCES = {"render_bwd": "render_bwd.cu", "render_fwd": "render_fwd.cu", "count_work": "render_fwd.cu",
           "preprocess_bwd": "preprocess_bwd.cu", "band_": "preprocess_bwd.cu", "preprocess_fwd": "preprocess.cu",
           "mark_visible": "preprocess.cu", "tile_": "binning.cu", "onesweep": "binning.cu", "radix_": "binning.cu",
           "rebuild_keys": "binning.cu", "exchange_": "exchange.cu", "ssim": "ssim.cu", "adam": "adam.cu",
           "densify": "densify.cu", "ply_": "densify.cu", "fp32_peak": "peak.cu", "loss_": "loss.cu"}


def src_sha1(kernel):
    for prefix, f in SOURCES.items():
        if kernel.startswith(prefix) and os.path.exists(os.path.join(CSRC, f)):
            return hashlib.sha1(open(os.path.join(CSRC, f), "rb").read()).hexdigest()[:16]
    return None

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}


def val(r, name, scale=1.0):
    if name not in col or r[col[name]] in ("", "n/a"):
        return None
    return float(r[col[name]].replace(",", "")) * scale


def to_unit(r, name, want):
    """time in us, bytes in MB regardless of the unit ncu picked for the column"""
    v = val(r, name)
    if v is None:
        return None
    u = units[col[name]]
    f = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u) if want == "us" else \
        {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u)
    return round(v * f, 1)


out = []
for r in rows[2:]:
    name = re.sub(r"^void ", "", r[col["Kernel Name"]])
    name = re.sub(r"\(.*$", "", name).replace("gsr::", "")
    name = name.replace("(bool)", "").replace("(int)", "")
    out.append(dict(kernel=name, us=to_unit(r, "gpu__time_duration.sum", "us"),
                    dram_read_MB=to_unit(r, "dram__bytes_read.sum", "MB"),
                    dram_write_MB=to_unit(r, "dram__bytes_write.sum", "MB"),
                    dram_pct=val(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
                    issue_active_pct=val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                    warps_active_pct=val(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
                    regs=int(val(r, "launch__registers_per_thread") or 0), inst=int(val(r, "smsp__inst_executed.sum") or 0),
                    # shared-memory wavefronts and L2 reduction / atomic sectors
                    smem_wavefronts=val(r, "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"),
                    smem_bank_conflicts=val(r, "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"),
                    l2_red_sectors=val(r, "lts__t_sectors_op_red.sum"), l2_atom_sectors=val(r, "lts__t_sectors_op_atom.sum"),
                    fma_pipe_pct=val(r, "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
                    alu_pipe_pct=val(r, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
                    nvlink_rx_MB=to_unit(r, "nvlrx__bytes.sum", "MB"), nvlink_tx_MB=to_unit(r, "nvltx__bytes.sum", "MB"),
                    src_sha1=src_sha1(name)))
print(json.dumps(out, indent=0))
