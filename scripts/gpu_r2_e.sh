#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2e_pytest.log
tail -4 gpurun_out/r2e_pytest.log
timeout 300 python scripts/err_check.py > gpurun_out/r2e_err.log 2>&1; cat gpurun_out/r2e_err.log | tail -8
timeout 600 python bench.py --steps 20 --warmup 3 --no-side-legs --no-cpu-baseline --profile-out gpurun_out/r2e_profile.csv > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2e_bench.json')); print(d['value'], d['ms_per_step'], d['roofline']['achieved'], d['roofline']['frac'], d['roofline']['generator_convs'])"
cat gpurun_out/r2e_profile.csv
