#!/bin/bash
# round-2 final evidence: tests, smoke, bench line with side records, reference arm, latency table
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2f_pytest.log; tail -3 gpurun_out/r2f_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2f_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/r2f_smoke.log
timeout 900 python bench.py --steps 20 --warmup 3 --profile-out gpurun_out/r2f_profile.csv > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo "bench rc=$?"
cat gpurun_out/r2f_bench.json; cat gpurun_out/r2f_profile.csv; tail -3 gpurun_out/r2f_bench.err
timeout 300 python scripts/gpu_latency.py > gpurun_out/r2f_latency.log 2>&1; cat gpurun_out/r2f_latency.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2f_ref.json 2> gpurun_out/r2f_ref.err; echo "ref rc=$?"; cat gpurun_out/r2f_ref.json
# ncu launch list of the same bench command (shares, not absolutes: cold-cache, serialised)
python bench.py --steps 2 --warmup 3 --no-side-legs --no-cpu-baseline > gpurun_out/r2f_ncu_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2f_ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-side-legs --no-cpu-baseline > gpurun_out/r2f_ncu_run.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/r2f_ncu_launches.csv
