#!/bin/bash
# GPU tests, A/B bench (product library vs ALT), then the aux kernels' ncu launch list
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
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"mrf_sum|cond_gemv|conv_post|pack" --csv --log-file gpurun_out/aux_launches.csv python scripts/one_call.py fp16 2 > gpurun_out/aux_ncu.log 2>&1
echo "ncu rc=$?"; grep "gpu__time_duration" gpurun_out/aux_launches.csv | awk -F'","' '{print $5, $NF}' | sed 's/void //; s/(const.*)//' | cut -c1-100
