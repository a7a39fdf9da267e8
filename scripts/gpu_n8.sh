#!/bin/bash
N=${N:-8}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 30 --warmup 3 > gpurun_out/m_bench_n$N.log 2>&1
echo "n$N rc=$?"; grep '^{' gpurun_out/m_bench_n$N.log | tail -1 | cut -c1-900; grep -i "error\|Traceback" gpurun_out/m_bench_n$N.log | head -5
