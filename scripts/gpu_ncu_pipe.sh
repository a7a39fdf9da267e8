#!/bin/bash
# ncu counters (tensor-pipe %, DRAM bytes, duration) for every launch of ONE C2 call (the second of two; 51 launches)
mkdir -p gpurun_out
M="gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_tensor.sum,sm__inst_executed_pipe_uniform.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_bytes.sum,launch__grid_size,launch__registers_per_thread"
python scripts/one_call.py fp16 2 > gpurun_out/ncu_pipe_plain.log 2>&1 &&
timeout 900 ncu --metrics $M --clock-control none --launch-skip 51 --launch-count 51 --csv --log-file gpurun_out/ncu_pipe.csv python scripts/one_call.py fp16 2 > gpurun_out/ncu_pipe_run.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_pipe_plain.log gpurun_out/ncu_pipe_run.log; wc -l gpurun_out/ncu_pipe.csv; head -5 gpurun_out/ncu_pipe.csv | cut -c1-400
