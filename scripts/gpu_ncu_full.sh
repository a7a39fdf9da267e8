#!/bin/bash
# ncu --set full of the launches of one C2 call (one_call.py turns graphs off so that each launch is a plain kernel); only the raw-page CSVs
# travel back (the .ncu-rep files are 70 MB each)
mkdir -p gpurun_out /tmp/ncu
timeout 900 ncu --set full --clock-control none -k regex:rb_fused --launch-skip 36 --launch-count 36 -f -o /tmp/ncu/rb_full python scripts/one_call.py fp16 2 > gpurun_out/j_ncu_full.log 2>&1
echo "rc=$?"; ncu -i /tmp/ncu/rb_full.ncu-rep --page raw --csv > gpurun_out/j_rb_full_raw.csv; ls -la gpurun_out/j_rb_full_raw.csv
if [ -n "$REST" ]; then
timeout 600 ncu --set full --clock-control none -k regex:"conv_tc|mrf_sum|conv_post|cond_gemv" --launch-skip 51 --launch-count 51 -f -o /tmp/ncu/rest_full python scripts/one_call.py fp16 2 > gpurun_out/j_ncu_rest.log 2>&1
echo "rc=$?"; ncu -i /tmp/ncu/rest_full.ncu-rep --page raw --csv > gpurun_out/j_rest_full_raw.csv; ls -la gpurun_out/j_rest_full_raw.csv
fi
