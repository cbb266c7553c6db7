#!/bin/bash
# One GPU-box visit: parity tests, factorization probe (both pivot variants), bench line.
mkdir -p gpurun_out
TAG=${1:-r2}
python -m pytest tests -m gpu -x -q -s 2>&1 | tail -80 > gpurun_out/${TAG}_tests.log
for n in 4096 8192 16384; do
  python tools/prof_potrf.py $n >> gpurun_out/${TAG}_potrf.log 2>&1
  HYB_POTF2_FAST=1 python tools/prof_potrf.py $n >> gpurun_out/${TAG}_potrf_fast.log 2>&1
done
HYB_POTRF_OB=512 python tools/prof_potrf.py 8192 >> gpurun_out/${TAG}_potrf_ob512.log 2>&1
python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
tail -5 gpurun_out/${TAG}_tests.log; cat gpurun_out/${TAG}_potrf.log gpurun_out/${TAG}_potrf_fast.log gpurun_out/${TAG}_potrf_ob512.log; head -c 600 gpurun_out/${TAG}_bench.json
