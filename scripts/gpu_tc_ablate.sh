#!/bin/bash
# conv_tc ablations with the debug-hook library (results are numerically wrong when a mask is set):
#   VITSB_TC_DBG: 1 = A loaded only once per slab-equivalent, 2 = no weight loads, 4 = no epilogue
#   OPTS: extra bench flags, e.g. OPTS="--option tc_pair=0"
mkdir -p gpurun_out
for m in ${MASKS:-0 4 2 1}; do
  VITSB_LIB=$PWD/vits_b200/libvitsb200_dbg.so VITSB_TC_DBG=$m timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-side-legs $OPTS --profile-out gpurun_out/tcabl_$m.csv > gpurun_out/tcabl_$m.log 2>&1
  echo "TC_DBG=$m $OPTS: $(grep -E 'ups|flow|conv_pre' gpurun_out/tcabl_$m.csv | sort | awk -F, '{printf "%s=%.3f ", $1, $2}')"
done
