#!/usr/bin/env python
"""Per-kernel SASS evidence from the built library (cuobjdump): registers, shared memory, and counts of the opcodes
that show how a kernel moves data and reduces (UBLKCP = cp.async.bulk through the TMA engine, SYNCS = mbarrier,
REDG / RED = reductions at L2, ATOM*, MUFU.* = special-function unit, SHFL, VOTE, LDS / STS widths, tensor-core
opcodes: none expected).  Usage: python profiles/sass_summary.py > profiles/r02/sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gaussian_splatting_b200", "libgsrast.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
res = subprocess.run(["cuobjdump", "-res-usage", so], capture_output=True, text=True).stdout
usage = {}
cur = None
for line in res.splitlines():
    m = re.search(r"Function (\S+):", line)
    if m:
        cur = m.group(1)
    m = re.search(r"REG:(\d+).*SHARED:(\d+)", line)
    if m and cur:
        usage[cur] = (int(m.group(1)), int(m.group(2)))
WATCH = ["UBLKCP", "SYNCS", "REDG", "RED.", "ATOMG", "ATOMS", "MUFU.EX2", "MUFU.RCP", "MUFU.RSQ", "MUFU.SQRT", "MUFU.LG2",
         "SHFL", "VOTE", "REDUX", "MATCH", "LDS.128", "LDS.64", "STS.128", "LDG.E.128", "STG.E.128", "HMMA", "IMMA", "UTCHMMA",
         "UTCQMMA", "BAR.SYNC", "FFMA", "LDL", "STL"]
counts = collections.OrderedDict()
name = None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = m.group(1)
        counts[name] = collections.Counter()
        continue
    if name is None:
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        counts[name]["_total"] += 1
        for w in WATCH:
            if op.startswith(w):
                counts[name][w] += 1
print(f"# {os.path.basename(so)}: SASS opcode counts per kernel (static instructions), registers / static shared bytes")
for fn, c in counts.items():
    dem = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip()
    dem = re.sub(r"\(.*$", "", dem).replace("gsr::", "")
    reg, smem = usage.get(fn, (0, 0))
    hits = " ".join(f"{w}={c[w]}" for w in WATCH if c[w])
    print(f"{dem:<44} regs={reg:<3} smem={smem:<6} sass={c['_total']:<5} {hits}")
