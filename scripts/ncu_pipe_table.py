"""Pivot the long-format ncu CSV of scripts/gpu_ncu_pipe.sh into one row per launch:
    python scripts/ncu_pipe_table.py gpurun_out/ncu_pipe.csv > profiles/r02_ncu_tensor_pipe.csv"""
import csv
import re
import sys

rows = {}
order = []
with open(sys.argv[1]) as f:
    rd = csv.reader(l for l in f if l.startswith('"'))
    hdr = next(rd)
    I = {h: i for i, h in enumerate(hdr)}
    for r in rd:
        k = r[I["ID"]]
        if k not in rows:
            rows[k] = {"kernel": r[I["Kernel Name"]], "grid": r[I["Grid Size"]], "block": r[I["Block Size"]]}
            order.append(k)
        rows[k][r[I["Metric Name"]]] = r[I["Metric Value"]].replace(",", "")


def short(name):
    m = re.search(r"rb_fused_kernel<(\w+), *(\d+), *(\d+)", name)
    if m:
        return f"rb_fused C={m.group(2)} k={m.group(3)}"
    m = re.search(r"(\w+_kernel)", name)
    return m.group(1) if m else name[:40]


cols = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "lts__t_sector_hit_rate.pct", "launch__registers_per_thread"]
print("launch,kernel,grid,time_us,tensor_pipe_pct_elapsed,tensor_pipe_pct_active,tensor_insts,sm_throughput_pct,dram_read_MB,dram_write_MB,l2_hit_pct,regs")
for n, k in enumerate(order):
    r = rows[k]
    g = lambda c: float(r.get(c, "nan") or "nan")
    print(f"{n},{short(r['kernel'])},{r['grid'].replace(',', ' ')},{g(cols[0]) / 1e3:.1f},{g(cols[1]):.1f},{g(cols[2]):.1f},{g(cols[3]):.0f},"
          f"{g(cols[4]):.1f},{g(cols[5]) / 1e6:.1f},{g(cols[6]) / 1e6:.1f},{g(cols[7]):.1f},{g(cols[8]):.0f}")
