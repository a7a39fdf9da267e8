#!/bin/bash
# A/B one environment switch of the library on the bench workload: ENVAB="NAME=a NAME=b"
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python scripts/err_check.py 2>&1 | tail -6
for kv in $ENVAB; do
  env $kv timeout 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline ${BENCHARGS} --profile-out gpurun_out/env_$kv.csv > gpurun_out/env_$kv.log 2>&1
  python - <<PY
import json
l=[x for x in open("gpurun_out/env_$kv.log") if x.startswith("{")]
if l:
    d=json.loads(l[-1]); print("$kv value", d["value"], "ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "e2e", d["e2e"]["value"], "launches", d["gpu_launches"])
else:
    print("$kv FAILED", open("gpurun_out/env_$kv.log").read()[-600:])
PY
  grep -E "flow" gpurun_out/env_$kv.csv | tr '\n' ' '; echo
done
