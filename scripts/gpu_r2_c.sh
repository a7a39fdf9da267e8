#!/bin/bash
# round 2: flow bring-up, full GPU test suite, bench line with side legs
mkdir -p gpurun_out
timeout 300 python scripts/gpu_flow_bringup.py > gpurun_out/r2c_bringup.log 2>&1; echo "bringup rc=$?"; tail -5 gpurun_out/r2c_bringup.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2c_pytest.log
tail -4 gpurun_out/r2c_pytest.log
timeout 600 python bench.py --steps 20 --warmup 3 --profile-out gpurun_out/r2c_profile.csv > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo "bench rc=$?"
cat gpurun_out/r2c_bench.json; cat gpurun_out/r2c_profile.csv; tail -5 gpurun_out/r2c_bench.err
