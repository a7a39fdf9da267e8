#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for ns in 6 8; do
VITSB_RB_NSLAB=$ns timeout 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline --profile-out gpurun_out/i_profile_ns$ns.csv > gpurun_out/i_bench_ns$ns.log 2>&1
python - <<PY
import json
l=[x for x in open("gpurun_out/i_bench_ns$ns.log") if x.startswith("{")]
d=json.loads(l[-1]); print("NSLAB=$ns value", d["value"], "ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "e2e", d["e2e"]["value"])
PY
grep "fused" gpurun_out/i_profile_ns$ns.csv
done
for t in 32,3 32,11 64,3 64,7 128,3 256,3; do
VITSB_LIB=$PWD/vits_b200/libvitsb200_dbg.so VITSB_RB_TRACE=$t,1,12 timeout 200 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/i_trace.log 2>&1
grep -A8 "rb_fused trace" gpurun_out/i_trace.log | cut -c1-700
done
