#!/bin/bash
# GPU tests, A/B bench (product library vs ALT), aux launch list timing, then ncu --set full of ups.2 / ups.3 (conv_tc launches 8, 9)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for rep in 1 2; do
for lib in libvitsb200.so ${ALT:-libvitsb200_prev.so}; do
VITSB_LIB=$PWD/vits_b200/$lib timeout 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline --no-side-legs --profile-out gpurun_out/ab_${lib}_$rep.csv > gpurun_out/ab_${lib}_$rep.log 2>&1
python - <<PY
import json
l=[x for x in open("gpurun_out/ab_${lib}_$rep.log") if x.startswith("{")]
d=json.loads(l[-1]); print("$lib rep $rep value", d["value"], "ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "e2e", d["e2e"]["value"])
PY
grep -E "post|gemv|mrf|ups" gpurun_out/ab_${lib}_$rep.csv | tr '\n' ' '; echo
done
done
timeout 300 python scripts/gpu_latency.py 2>&1 | tail -7
K=conv_tc SKIP=8 COUNT=2 OUT=ups23 bash scripts/gpu_ncu_kernel.sh
