#!/bin/bash
# end-of-round evidence (state i): tests, smoke, bench line + family table, reference arm, latency table,
# ncu launch list of the bench command, ncu counter table of every launch of one C2 call
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/i_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/i_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/i_smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py --profile-out gpurun_out/i_profile.csv > gpurun_out/i_bench.json 2> gpurun_out/i_bench.err; echo "bench rc=$?"
timeout 300 python scripts/gpu_latency.py > gpurun_out/i_latency.log 2>&1; echo "latency rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/i_ref.json 2> gpurun_out/i_ref.err; echo "ref rc=$?"
python bench.py --steps 2 --warmup 3 --no-side-legs --no-cpu-baseline > gpurun_out/i_ncu_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/i_ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-side-legs --no-cpu-baseline > gpurun_out/i_ncu_run.log 2>&1
echo "ncu launches rc=$?"; wc -l gpurun_out/i_ncu_launches.csv
bash scripts/gpu_ncu_pipe.sh > gpurun_out/i_ncu_pipe.log 2>&1; echo "ncu pipe rc=$?"
