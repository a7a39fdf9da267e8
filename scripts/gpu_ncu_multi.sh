#!/bin/bash
# ncu --set full with source-level warp sampling for several single launches of one C2 call: SPECS="name:kernel_regex:skip ..."
mkdir -p gpurun_out
SPECS=${SPECS:-c128k3:rb_fused:45 c32k3:rb_fused:63 c64k7:rb_fused:57}
for spec in $SPECS; do
  IFS=: read name k skip <<< "$spec"
  timeout 600 ncu --set full --warp-sampling-interval 0 --import-source on --clock-control none -k regex:$k --launch-skip $skip --launch-count 1 -f -o gpurun_out/one_$name python scripts/one_call.py fp16 2 > gpurun_out/one_${name}.log 2>&1
  echo "$name rc=$?"
done
ls -la gpurun_out/*.ncu-rep
