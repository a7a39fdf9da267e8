#!/usr/bin/env python
"""Warp-state samples of one profiled launch, by code region and stall reason.

    python scripts/ncu_stalls.py gpurun_out/x.ncu-rep [kernel_index] [lo-hi ...]

Without ranges: totals per stall reason and the 40 hottest SASS lines with their dominant reason.  With SASS line ranges
(e.g. 900-1300 1300-2200): the same per range (a warp role's loop).
"""
import csv, io, subprocess, sys
rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
ranges = [tuple(int(x) for x in a.split("-")) for a in sys.argv[3:]]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
ks, cur = [], None
for r in csv.reader(io.StringIO(src)):
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}; ks.append(cur); continue
    if cur is not None: cur["rows"].append(r)
k = ks[which]
h, body = k["rows"][0], k["rows"][1:]
si, so, ie = h.index("# Samples"), h.index("Source"), h.index("Instructions Executed")
reasons = [(i, c) for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
print(k["name"][:90], "| SASS lines", len(body), "| samples", sum(int(r[si] or 0) for r in body))
def report(lo, hi, topn):
    rows = body[lo:hi]
    tot = sum(int(r[si] or 0) for r in rows)
    agg = sorted(((sum(int(r[i] or 0) for r in rows), c) for i, c in reasons), reverse=True)
    print(f"--- lines {lo}-{hi}: {tot} samples | " + " ".join(f"{c[6:]}={v}" for v, c in agg if v > tot * 0.01))
    top = sorted(range(len(rows)), key=lambda i: -int(rows[i][si] or 0))[:topn]
    for i in sorted(top):
        r = rows[i]
        dom = max(reasons, key=lambda ic: int(r[ic[0]] or 0))[1][6:]
        prev = rows[i - 1][so].strip()[:60] if i > 0 else ""
        print(f"{lo + i:5d} {r[si]:>6} x{r[ie]:>8} {dom:<12} {r[so].strip()[:70]:<70} <- {prev}")
if ranges:
    for lo, hi in ranges: report(lo, hi, 14)
else:
    report(0, len(body), 40)
