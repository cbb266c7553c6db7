#!/bin/bash
# bench.py at G GPUs: N = 8192 in both multi-GPU forms, N = 16384 in the single-process form
G=${1:-2}; TAG=${2:-r2s}
mkdir -p gpurun_out
run() {  # mode N steps
  local mode=$1 N=$2 steps=$3
  HYB_MULTI_MODE=$mode python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $G --steps $steps --warmup 3 --N $N > gpurun_out/${TAG}_g${G}_N${N}_${mode}.json 2> gpurun_out/${TAG}_g${G}_N${N}_${mode}.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}_g${G}_N${N}_${mode}.json"))
    print("G=$G N=$N ${mode}: ms/step", round(d["ms_per_step"],2), "e2e it/s", round(d["e2e"]["value"],2), "asm", round(d["schur_assemble_ms"],2), "reduce", round(d["schur_reduce_ms"],2), "chol", round(d["cholesky_ms"],2), {k: round(v,2) for k,v in d["phase_ms"].items()})
except Exception as e:
    print("G=$G N=$N ${mode} failed", e)
PY
}
run single_process 8192 20
run rootworker 8192 20
run single_process 16384 10
