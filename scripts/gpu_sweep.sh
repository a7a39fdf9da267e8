#!/bin/bash
# C4 sweep points on one GPU: utterances per step 16 (C2 headline) / 64 / 128
mkdir -p gpurun_out
for b in 16 64 128; do
  timeout 600 python bench.py --batch-per-gpu $b --steps 10 --warmup 3 --no-cpu-baseline --profile-out gpurun_out/sweep_b$b.csv > gpurun_out/sweep_b$b.log 2>&1
  python - <<PY
import json
l=[x for x in open("gpurun_out/sweep_b$b.log") if x.startswith("{")]
if l:
    d=json.loads(l[-1]); print("batch $b: value %.0f audio-s/s, %.2f ms/step, audio/step %.1f s, frac %.3f, e2e %.0f" % (d["value"], d["ms_per_step"], d["config"]["valid_audio_s_per_step"], d["roofline"]["frac"], d["e2e"]["value"]))
else:
    print("batch $b failed:", open("gpurun_out/sweep_b$b.log").read()[-400:])
PY
  grep -E "flow|mrf|fused" gpurun_out/sweep_b$b.csv | tr '\n' ' '; echo
done
