#!/bin/bash
# syrk rasterisation: timing and ncu --set full DRAM / L2 counters, plain triangle walk (SB=0) vs super-blocks
mkdir -p gpurun_out
TAG=${1:-r2m}
python -m pytest tests/test_gpu_dense.py tests/test_gpu_schur.py -m gpu -x -q 2>&1 | tail -3
for sb in 0 8 12 16; do echo "SB=$sb"; HYB_GEMM_SB=$sb python tools/prof_gemm.py | tail -1; done
python tools/prof_gemm.py > gpurun_out/${TAG}_plain.log 2>&1 || exit 1
for sb in 0 12; do
  HYB_GEMM_SB=$sb ncu --set full --clock-control none -k regex:gemm_tn_kernel -s 2 -c 1 -f -o gpurun_out/${TAG}_syrk_sb$sb python tools/prof_gemm.py > gpurun_out/${TAG}_syrk_sb$sb.log 2>&1
  ncu -i gpurun_out/${TAG}_syrk_sb$sb.ncu-rep --page raw --csv > gpurun_out/${TAG}_syrk_sb${sb}_raw.csv 2>/dev/null
  rm -f gpurun_out/${TAG}_syrk_sb$sb.ncu-rep
done
python tools/ncu_summary2.py gpurun_out/${TAG}_syrk_sb0_raw.csv gpurun_out/${TAG}_syrk_sb12_raw.csv | grep -E "file|duration_us|dram_bytes|l2_hit|bank|dmma|tensor"
