#!/bin/bash
# GPU tests, A/B bench (product library vs ALT), then ncu --set full of the kernels matching $K of the second C2 call
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
grep -E "post|gemv|mrf|ups|conv_pre" gpurun_out/ab_${lib}_$rep.csv | tr '\n' ' '; echo
done
done
K=${K:-"conv_post_rows|cond_gemv_regs"} SKIP=${SKIP:-2} COUNT=${COUNT:-2} OUT=${OUT:-auxk} bash scripts/gpu_ncu_kernel.sh
