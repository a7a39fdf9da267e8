#!/bin/bash
# ncu --set full with source-level sampling for ONE rb_fused launch (default: first C=64 k=3 launch of the second call)
mkdir -p gpurun_out
SKIP=${SKIP:-54}   # 36 launches of call 1 + 18 (stages 0,1) of call 2 -> first launch of stage 2 (C=64, k=3)
timeout 600 ncu --set full --warp-sampling-interval 0 --import-source on --clock-control none -k regex:rb_fused --launch-skip $SKIP --launch-count 1 -f -o gpurun_out/one_rb python scripts/one_call.py fp16 2 > gpurun_out/one_ncu.log 2>&1
echo "rc=$?"; ls -la gpurun_out/one_rb.ncu-rep
