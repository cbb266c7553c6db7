#!/bin/bash
mkdir -p gpurun_out
TAG=${1:-r2c}
python -m pytest tests -m gpu -x -q -s 2>&1 | tail -150 > gpurun_out/${TAG}_tests.log
tail -5 gpurun_out/${TAG}_tests.log
HYB_POTF2_FAST=1 python -m pytest tests/test_gpu_schur.py tests/test_golden.py tests/test_gpu_dense.py -m gpu -q 2>&1 | tail -15 > gpurun_out/${TAG}_tests_fast.log
tail -5 gpurun_out/${TAG}_tests_fast.log
for n in 4096 8192 16384; do
  python tools/prof_potrf.py $n 2>&1 | tail -1
  HYB_POTF2_FAST=1 python tools/prof_potrf.py $n 2>&1 | tail -1
done
HYB_POTRF_TRACE=1 python tools/prof_potrf.py 8192 2>&1 | tail -36 > gpurun_out/${TAG}_trace.txt
