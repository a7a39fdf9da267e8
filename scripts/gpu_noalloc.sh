#!/bin/bash
mkdir -p gpurun_out
# alone, in a fresh process
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -15 gpurun_out/noalloc_alone.log
# after all the other GPU tests
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/noalloc_all.log 2>&1; echo "all rc=$?"; tail -6 gpurun_out/noalloc_all.log
