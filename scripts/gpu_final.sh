#!/bin/bash
# Round-end evidence: GPU tests, default bench (with CPU baseline), reference arm, ncu launch list of the same bench command
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py --profile-out gpurun_out/z_profile.csv > gpurun_out/z_bench.log 2>&1; echo "bench rc=$?"
grep '^{' gpurun_out/z_bench.log | tail -1
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/z_ref.log 2>&1; echo "ref rc=$?"
grep '^{' gpurun_out/z_ref.log | tail -1 | cut -c1-400
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/z_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/z_ncu.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/z_launches.csv
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -4
