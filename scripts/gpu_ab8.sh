#!/bin/bash
# bench A/B of several builds (per-family event timings): LIBS="a.so b.so ..."
mkdir -p gpurun_out
for rep in 1 2; do
for lib in ${LIBS:-libvitsb200.so libvitsb200_alt.so libvitsb200_prev.so}; do
  VITSB_LIB=$PWD/vits_b200/$lib timeout 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline --no-side-legs --profile-out gpurun_out/ab_$lib.csv > gpurun_out/ab_$lib.log 2>&1
  python - <<PY
import json
l=[x for x in open("gpurun_out/ab_$lib.log") if x.startswith("{")]
d=json.loads(l[-1]); print("$lib value %.0f ms %.3f frac %.3f e2e %.0f" % (d["value"], d["ms_per_step"], d["roofline"]["frac"], d["e2e"]["value"]))
PY
  grep -E "ups|conv_pre" gpurun_out/ab_$lib.csv | sort | tr '\n' ' '; echo
done
done
