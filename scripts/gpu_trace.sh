#!/bin/bash
# CTA-0 timelines of the fused ResBlock kernel (debug-hook library) + full ncu launch list of the product library
mkdir -p gpurun_out
for t in ${TRACES:-"32,3 32,11 64,3 64,7 64,11"}; do
  VITSB_LIB=$PWD/vits_b200/libvitsb200_dbg.so VITSB_RB_TRACE=$t,1,24 timeout 200 python bench.py --steps 1 --warmup 3 --no-cpu-baseline 2>&1 | grep -A8 "rb_fused trace" | cut -c1-400
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/g_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/g_ncu.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/g_launches.csv
