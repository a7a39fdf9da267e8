#!/bin/bash
# ncu --set full of ONE launch of the kernel(s) matching $K (regex) in one C2 call; the report comes back in gpurun_out/
K=${K:-cond_gemv}; SKIP=${SKIP:-1}; OUT=${OUT:-one_kernel}
mkdir -p gpurun_out
timeout 300 python scripts/one_call.py fp16 2 > gpurun_out/${OUT}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${OUT}_plain.log; exit 1; }
timeout 600 ncu --set full --import-source on --clock-control none -k regex:"$K" --launch-skip $SKIP --launch-count ${COUNT:-1} -f -o gpurun_out/$OUT python scripts/one_call.py fp16 2 > gpurun_out/${OUT}_ncu.log 2>&1
echo "rc=$?"; ls -la gpurun_out/$OUT.ncu-rep
