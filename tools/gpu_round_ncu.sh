#!/bin/bash
# ncu --set full captures of the HBM-bound kernels (and the Cholesky panel kernels) inside real runs.
mkdir -p gpurun_out
TAG=${1:-r2g}
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
CFG4="python tools/run_cfg.py cfg4s"
CFG4N="python tools/run_cfg.py cfg4ns"
$BENCH > gpurun_out/${TAG}_plain_bench.log 2>&1 || { echo "plain bench failed"; tail -5 gpurun_out/${TAG}_plain_bench.log; exit 1; }
$CFG4 > gpurun_out/${TAG}_plain_cfg4.log 2>&1 || { echo "plain cfg4 failed"; tail -5 gpurun_out/${TAG}_plain_cfg4.log; exit 1; }
$CFG4N > gpurun_out/${TAG}_plain_cfg4n.log 2>&1 || { echo "plain cfg4n failed"; tail -5 gpurun_out/${TAG}_plain_cfg4n.log; exit 1; }
cap() {  # name regex skip count cmd...
  local name=$1 rx=$2 skip=$3 cnt=$4; shift 4
  ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c $cnt -f -o gpurun_out/${TAG}_ncu_$name "$@" > gpurun_out/${TAG}_ncu_$name.log 2>&1
  echo "$name rc=$?"
  # gpurun_out/ may carry 64 MiB at most: keep the raw-counter page as CSV, drop the report (two are kept whole)
  ncu -i gpurun_out/${TAG}_ncu_$name.ncu-rep --page raw --csv > gpurun_out/${TAG}_ncu_${name}_raw.csv 2>/dev/null
  case $name in gemm_t|potf2) ;; *) rm -f gpurun_out/${TAG}_ncu_$name.ncu-rep ;; esac
}
cap gemm_t        'gemm_t_kernel'         6 2 $BENCH
cap gemm_n        'gemm_n_partial_kernel' 6 2 $BENCH
cap trsv_fwd      'trsv_fwd_kernel'       6 2 $BENCH
cap trsv_bwd      'trsv_bwd_kernel'       6 2 $BENCH
cap potf2         'potf2_kernel'         70 2 $BENCH
cap trsm          'trsm_upper_T_kernel'  70 2 $BENCH
cap psd_update    'psd_update_kernel'     6 2 $BENCH
cap soc_prod_wide 'soc_prod_kernel'       0 2 $BENCH
cap cand          'cand_kernel'           6 2 $BENCH
cap qr_apply      'qr_apply_kernel'    2000 2 $BENCH
cap soc_update4   'soc_update_kernel'    10 2 $CFG4
cap soc_prod4     'soc_prod_kernel'      10 3 $CFG4
cap ninf_update4  'ninf_update_kernel'   10 2 $CFG4N
cap ninf_prod4    'ninf_prod_kernel'     10 3 $CFG4N
cap nonneg_prod4  'nonneg_prod_kernel'   10 2 $CFG4
ls gpurun_out/${TAG}_ncu_*_raw.csv | wc -l; du -sh gpurun_out
