#!/bin/bash
# one `ncu --set full` capture (kernel regex $1, skip $2, count $3) inside the bench command, after it exited 0 plain
mkdir -p gpurun_out
RX=$1; SKIP=$2; CNT=$3; NAME=$4
CMD="python tools/prof_potrf.py 8192"
$CMD > gpurun_out/ncu_one_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_one_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:$RX -s $SKIP -c $CNT -f -o gpurun_out/r02_ncu_$NAME $CMD > gpurun_out/ncu_one.log 2>&1
echo "ncu rc=$?"
ncu -i gpurun_out/r02_ncu_$NAME.ncu-rep --page raw --csv > gpurun_out/r02_ncu_${NAME}_raw.csv 2>/dev/null
ncu -i gpurun_out/r02_ncu_$NAME.ncu-rep --page source --csv > gpurun_out/r02_ncu_${NAME}_source.csv 2>/dev/null
rm -f gpurun_out/r02_ncu_$NAME.ncu-rep
du -sh gpurun_out; tail -3 gpurun_out/ncu_one.log
