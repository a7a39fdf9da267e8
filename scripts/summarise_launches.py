#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name.

    python scripts/summarise_launches.py gpurun_out/x_launches.csv > profiles/rNN_x_launch_summary.csv
"""
import csv
import re
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(open(sys.argv[1])) if r and r[0].isdigit()]
agg = OrderedDict()
for r in rows:
    name = re.sub(r"\(.*", "", r[4]).replace("void ", "").strip()
    ns = float(r[-1])
    a = agg.setdefault(name, [0, 0.0, r[7], r[8]])
    a[0] += 1
    a[1] += ns
tot = sum(a[1] for a in agg.values())
print("kernel,launches,total_us,avg_us,share,block,grid_of_first")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"\"{k}\",{a[0]},{a[1] / 1e3:.1f},{a[1] / 1e3 / a[0]:.2f},{a[1] / tot:.4f},\"{a[2]}\",\"{a[3]}\"")
print(f"TOTAL,{len(rows)},{tot / 1e3:.1f},,1.0,,")
