#!/usr/bin/env python
"""`ncu -i X.ncu-rep --page raw --csv` -> compact per-launch JSON (what bench.py's ncu_evidence() reads).
Usage: python profiles/ncu_kernels_json.py raw.csv > profiles/r01g_kernels.json"""
import csv
import json
import re
import sys

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
                    regs=int(val(r, "launch__registers_per_thread") or 0), inst=int(val(r, "smsp__inst_executed.sum") or 0)))
print(json.dumps(out, indent=0))
