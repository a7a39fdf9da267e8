#!/usr/bin/env python
"""profiles/traffic.json from the per-launch ncu counter table (scripts/ncu_pipe_table.py output):
    python scripts/traffic_from_pipe.py profiles/r02_i_ncu_tensor_pipe.csv > profiles/traffic.json"""
import csv, json, sys
rows = list(csv.DictReader(open(sys.argv[1])))
tens = [r for r in rows if float(r["tensor_insts"]) > 0]
gen = [r for r in tens if "flow" not in r["kernel"]]
t = lambda rs: sum(float(r["time_us"]) for r in rs)
wp = lambda rs: sum(float(r["time_us"]) * float(r["tensor_pipe_pct_elapsed"]) for r in rs) / t(rs)
dram = lambda rs: sum(float(r["dram_read_MB"]) + float(r["dram_write_MB"]) for r in rs) * 1e6
share = {}
for r in rows:
    k = "rb_fused" if r["kernel"].startswith("rb_fused") else r["kernel"]
    share[k] = share.get(k, 0.0) + float(r["time_us"])
tot = sum(share.values())
print(json.dumps({
    "source": f"{sys.argv[1]} (ncu counters, one C2 call, graphs off, cold L2 per launch)",
    "tensor_launches_per_step": len(tens),
    "dram_bytes_per_launch": dram(tens) / len(tens),
    "dram_bytes_per_step_all_launches": dram(rows),
    "algorithmic_note": "36 fused ResBlock launches move 2 B in + 2 B out per element = 5.9 GB; upsamplers ~0.5 GB; flow 4 x 14.6 MB (DRAM writes below the algorithmic figure: part of every output stream is still in L2 when ncu stops counting)",
    "generator_convs_tensor_pipe_pct_time_weighted": wp(gen),
    "tensor_launches_tensor_pipe_pct_time_weighted": wp(tens),
    "sum_launch_time_us": tot,
    "share_of_device_time": {k: round(v / tot, 4) for k, v in share.items()},
}, indent=1))
