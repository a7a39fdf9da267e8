#!/bin/bash
# A/B two builds of the library on the bench workload: product (libvitsb200.so) vs VITSB_LIB_ALT
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python scripts/err_check.py 2>&1 | tail -6
for lib in libvitsb200.so ${ALT:-libvitsb200_alt.so}; do
VITSB_LIB=$PWD/vits_b200/$lib timeout 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline --profile-out gpurun_out/ab_$lib.csv > gpurun_out/ab_$lib.log 2>&1
python - <<PY
import json
l=[x for x in open("gpurun_out/ab_$lib.log") if x.startswith("{")]
d=json.loads(l[-1]); print("$lib value", d["value"], "ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "e2e", d["e2e"]["value"])
PY
grep -E "fused|post|gemv|mrf" gpurun_out/ab_$lib.csv | tr '\n' ' '; echo
done
