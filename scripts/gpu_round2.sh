#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 300 python scripts/err_check.py 2>&1 | tail -6
for m in 1 0; do
VITSB_MRF_SEP=$m timeout 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline --profile-out gpurun_out/h_profile_sep$m.csv > gpurun_out/h_bench_sep$m.log 2>&1
python - <<PY
import json
l=[x for x in open("gpurun_out/h_bench_sep$m.log") if x.startswith("{")]
d=json.loads(l[-1]); print("MRF_SEP=$m value", d["value"], "ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "e2e", d["e2e"]["value"])
PY
done
cat gpurun_out/h_profile_sep1.csv
VITSB_LIB=$PWD/vits_b200/libvitsb200_dbg.so VITSB_RB_TRACE=32,3,1,24 timeout 200 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/h_trace.log 2>&1
grep -c . gpurun_out/h_trace.log; grep -A8 "rb_fused trace" gpurun_out/h_trace.log | cut -c1-600
