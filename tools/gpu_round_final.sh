#!/bin/bash
# Final validation visit on one GPU: whole GPU suite, smoke(), the bench line and the reference arm.
# (HYB_FINAL_ALL=1 adds the factorization sizes, the update-form GEMM rates and every BASELINE config.)
mkdir -p gpurun_out
TAG=${1:-r02_final}
python -m pytest tests -m gpu -q -s 2>&1 | tail -150 > gpurun_out/${TAG}_tests.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${TAG}_smoke.log 2>&1
if [ -n "$HYB_FINAL_ALL" ]; then
  for n in 465 1891 4096 8192 16384; do python tools/prof_potrf.py $n >> gpurun_out/${TAG}_potrf.log 2>&1; done
  HYB_POTRF_TRACE=1 python tools/prof_potrf.py 8192 >> gpurun_out/${TAG}_potrf_trace.log 2>&1
  python tools/prof_gemm_k.py > gpurun_out/${TAG}_gemm_k.log 2>&1
  for c in cfg1 cfg2 cfg3 cfg4s cfg4ns cfg4; do python tools/run_cfg.py $c > gpurun_out/${TAG}_run_$c.log 2>&1; done
fi
python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
tail -3 gpurun_out/${TAG}_tests.log; tail -2 gpurun_out/${TAG}_smoke.log; head -c 400 gpurun_out/${TAG}_bench.json; echo; head -c 300 gpurun_out/${TAG}_bench_ref.json
