#!/bin/bash
# round 2, first GPU call: fused-flow bring-up, full GPU test suite, one bench line
mkdir -p gpurun_out
timeout 300 python scripts/gpu_flow_bringup.py > gpurun_out/r2a_bringup.log 2>&1; echo "bringup rc=$?" | tee -a gpurun_out/r2a_bringup.log
tail -15 gpurun_out/r2a_bringup.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2a_pytest.log
tail -25 gpurun_out/r2a_pytest.log
timeout 300 python bench.py --steps 20 --warmup 3 --profile-out gpurun_out/r2a_profile.csv > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"
cat gpurun_out/r2a_bench.json; cat gpurun_out/r2a_profile.csv; tail -5 gpurun_out/r2a_bench.err
