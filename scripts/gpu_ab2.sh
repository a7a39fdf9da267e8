#!/bin/bash
# interleaved A/B of differently built libraries on the bench workload: LIBS="a.so b.so" ROUNDS=2 bash scripts/gpu_ab2.sh
mkdir -p gpurun_out
LIBS=${LIBS:-"libvitsb200_base.so libvitsb200.so"}
for r in $(seq 1 ${ROUNDS:-2}); do
for lib in $LIBS; do
VITSB_LIB=$PWD/vits_b200/$lib timeout 300 python bench.py --steps 40 --warmup 3 --no-cpu-baseline --no-side-legs --profile-out gpurun_out/ab_${lib}_$r.csv > gpurun_out/ab_${lib}_$r.log 2>&1
python - <<PY
import json
l=[x for x in open("gpurun_out/ab_${lib}_$r.log") if x.startswith("{")]
d=json.loads(l[-1]); print("$lib r$r value %.0f ms %.4f e2e %.0f" % (d["value"], d["ms_per_step"], d["e2e"]["value"]))
PY
awk -F, 'NR>2{printf "%s=%s ", $1, $2}' gpurun_out/ab_${lib}_$r.csv; echo
done
done
if [ -n "${TRACE_SHAPES:-64,3 64,7 32,3 128,3}" ] && [ -f vits_b200/libvitsb200_dbg.so ]; then
for shape in ${TRACE_SHAPES:-64,3 64,7 32,3 128,3}; do
  VITSB_LIB=vits_b200/libvitsb200_dbg.so VITSB_RB_TRACE="$shape,1,6" timeout 120 python scripts/one_call.py fp16 1 2>&1 | grep -A8 "rb_fused trace"
done > gpurun_out/rb_traces.txt
cat gpurun_out/rb_traces.txt
fi
