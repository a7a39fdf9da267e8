#!/bin/bash
# CTA-0 timelines of conv_tc_kernel (debug-hook library) for the shapes in $TRACES ("Cin,N" pairs), items 8.. of CTA 0's sequence
mkdir -p gpurun_out
for t in ${TRACES:-"64,64 128,128 512,2048"}; do
  VITSB_LIB=$PWD/vits_b200/libvitsb200_dbg.so VITSB_TC_TRACE=$t,1,${FIRST:-8} timeout 200 python scripts/one_call.py fp16 1 2>&1 | grep -A6 "conv_tc trace" | cut -c1-330
done
