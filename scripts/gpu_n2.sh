#!/bin/bash
# 2-GPU bench under torchrun (as the driver launches it) + reference arm + 1-GPU default bench
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 30 --warmup 3 > gpurun_out/k_bench_n2.log 2>&1
echo "n2 rc=$?"; grep '^{' gpurun_out/k_bench_n2.log | tail -1 | cut -c1-700
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/k_ref_n2.log 2>&1
echo "ref n2 rc=$?"; grep '^{' gpurun_out/k_ref_n2.log | tail -1 | cut -c1-600
timeout 600 python bench.py > gpurun_out/k_bench_n1.log 2>&1
echo "n1 rc=$?"; grep '^{' gpurun_out/k_bench_n1.log | tail -1
