#!/bin/bash
# ncu launch list (gpu__time_duration per launch) of the bench command, after the same command exited 0 plain
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/ncu_list_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_list_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 16450 -c 4000 --csv --log-file gpurun_out/r02_bench_N8192_launches_ncu.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/r02_bench_N8192_launches_ncu.csv; du -sh gpurun_out
