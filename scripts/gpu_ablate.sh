#!/bin/bash
# Timing ablations of the fused ResBlock kernel (results are numerically wrong when a mask is set):
#   1 no output stores, 2 no residual / acc loads, 4 no gate math / G stores, 8 no MMAs, 32 no TMEM loads in the store epilogue,
#   64 no slab TMA loads, 128 no weight-ring TMA loads
for m in ${MASKS:-0 3 47 111 175 239 64 128 192 8}; do
  VITSB_LIB=$PWD/vits_b200/libvitsb200_dbg.so VITSB_RB_DBG=$m timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --profile-out gpurun_out/abl_$m.csv > gpurun_out/abl_$m.log 2>&1
  echo "mask $m: $(grep resblock gpurun_out/abl_$m.csv | sort | awk -F, '{printf "%s=%.3f ", $1, $2}')"
done
