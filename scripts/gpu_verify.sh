#!/bin/bash
# head-of-tree check: GPU tests, smoke, the default bench line (what the driver runs at round end)
mkdir -p gpurun_out
S=$(date +%s)
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/v_pytest.log 2>&1; echo "pytest rc=$? $(( $(date +%s)-S ))s"; tail -3 gpurun_out/v_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/v_smoke.log 2>&1; echo "smoke rc=$? $(( $(date +%s)-S ))s"; tail -2 gpurun_out/v_smoke.log
timeout 900 python bench.py --profile-out gpurun_out/v_profile.csv > gpurun_out/v_bench.json 2> gpurun_out/v_bench.err; echo "bench rc=$? $(( $(date +%s)-S ))s"
cat gpurun_out/v_bench.json; cat gpurun_out/v_profile.csv
