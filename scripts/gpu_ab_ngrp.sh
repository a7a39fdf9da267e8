#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/ab_pytest.log 2>&1; echo "pytest(NGRP=3) rc=$?"; tail -2 gpurun_out/ab_pytest.log
for lib in libvitsb200.so libvitsb200_g2.so libvitsb200.so libvitsb200_g2.so; do
  VITSB_LIB=vits_b200/$lib timeout 600 python bench.py --steps 20 --warmup 3 --no-side-legs --no-cpu-baseline --profile-out gpurun_out/ab_$lib.csv > gpurun_out/ab_$lib.json 2>/dev/null
  python -c "
import json,sys; d=json.load(open('gpurun_out/ab_$lib.json')); print('$lib', round(d['value']), round(d['ms_per_step'],3), d['roofline']['generator_convs']['frac_burst'])"
  grep "resblock" gpurun_out/ab_$lib.csv
done
