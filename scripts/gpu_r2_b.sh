#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2b_pytest.log
tail -25 gpurun_out/r2b_pytest.log
VITSB_LIB=vits_b200/libvitsb200_dbg.so VITSB_FF_TRACE=1 timeout 120 python scripts/gpu_flow_trace.py c2 > gpurun_out/r2b_trace_c2.log 2>&1
VITSB_LIB=vits_b200/libvitsb200_dbg.so VITSB_FF_TRACE=1 timeout 120 python scripts/gpu_flow_trace.py one > gpurun_out/r2b_trace_one.log 2>&1
cat gpurun_out/r2b_trace_c2.log gpurun_out/r2b_trace_one.log
