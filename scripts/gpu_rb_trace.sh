#!/bin/bash
# clock64 timelines of CTA 0 of the fused ResBlock kernel (debug-hook library), tiles 8.. of its sequence
mkdir -p gpurun_out
for shape in "128,3" "256,3" "64,3" "128,11"; do
  VITSB_LIB=vits_b200/libvitsb200_dbg.so VITSB_RB_TRACE="$shape,1,6" timeout 120 python scripts/one_call.py fp16 1 2>&1 | grep -A8 "rb_fused trace"
done > gpurun_out/rb_traces.txt
cat gpurun_out/rb_traces.txt
