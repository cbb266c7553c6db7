#!/bin/bash
mkdir -p gpurun_out
TAG=${1:-r2b}
HYB_POTRF_TRACE=1 python tools/prof_potrf.py 8192 > gpurun_out/${TAG}_trace.txt 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/${TAG}_potrf_launches.csv python tools/prof_potrf.py 8192 > gpurun_out/${TAG}_ncu.log 2>&1
python -m pytest tests -m gpu -x -q -s 2>&1 | tail -120 > gpurun_out/${TAG}_tests.log
tail -5 gpurun_out/${TAG}_tests.log
