#!/bin/bash
# One GPU-box pass: error report + bench for both in_layer precisions, then the ncu launch list of the default build.
mkdir -p gpurun_out
for s in 1 0; do
  echo "== VITSB_FLOW_IN_SPLIT=$s"
  VITSB_FLOW_IN_SPLIT=$s timeout 300 python scripts/err_check.py 2>&1 | tail -7
  VITSB_FLOW_IN_SPLIT=$s timeout 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline --profile-out gpurun_out/g_profile_split$s.csv > gpurun_out/g_bench_split$s.log 2>&1
  python - <<PY
import json
l=[x for x in open("gpurun_out/g_bench_split$s.log") if x.startswith("{")]
d=json.loads(l[-1]); print("value", d["value"], "ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "e2e", d["e2e"]["value"])
PY
  grep flow gpurun_out/g_profile_split$s.csv
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/g_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/g_ncu.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/g_launches.csv
