#!/usr/bin/env python
"""Read an .ncu-rep: per-launch headline metrics and the hottest SASS lines per launch.

    python scripts/ncu_read.py gpurun_out/x.ncu-rep [launch_index] [top_n]
"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else None
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 25

raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
keys = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "lts__t_bytes.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct"]
for i, h in enumerate(hdr):
    if any(h == k or h.endswith("." + k) or h.endswith(k) for k in keys):
        print(f"{h} [{units[i]}]:", " | ".join(r[i][-40:] for r in data))

if which is not None:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True).stdout
    ks, cur = [], None
    for r in csv.reader(io.StringIO(src)):
        if r and r[0] == "Kernel Name":
            cur = []; ks.append(cur); continue
        if cur is not None:
            cur.append(r)
    k = ks[which * (len(ks) // len(data))]
    h, body = k[0], k[1:]
    si, so, ie = h.index("# Samples"), h.index("Source"), h.index("Instructions Executed")
    tot = sum(int(r[si] or 0) for r in body)
    print(f"launch {which}: {tot} samples, {len(body)} SASS lines")
    top = sorted(range(len(body)), key=lambda i: -int(body[i][si] or 0))[:topn]
    for i in sorted(top):
        ctx = ""
        if "BRA" in body[i][so] and i > 0:
            ctx = "   <- " + body[i - 1][so][:70]
        print(f"{i:5d} {body[i][si]:>6} {body[i][ie]:>9}  {body[i][so][:80]}{ctx}")
