#!/bin/bash
mkdir -p gpurun_out
TAG=${1:-r2f}
nvidia-smi -L > gpurun_out/${TAG}_gpus.txt
python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/${TAG}_multi_tests.log
tail -3 gpurun_out/${TAG}_multi_tests.log
for mode in rootworker single_process; do
  HYB_MULTI_MODE=$mode python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/${TAG}_bench2_${mode}.json 2> gpurun_out/${TAG}_bench2_${mode}.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}_bench2_${mode}.json"))
    print("${mode}", d["n_gpus"], "ms/step", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["value"],2), "asm", round(d["schur_assemble_ms"],2), "reduce", round(d["schur_reduce_ms"],2), "chol", round(d["cholesky_ms"],2), d["phase_ms"])
except Exception as e:
    print("${mode} failed", e)
PY
done
tail -5 gpurun_out/${TAG}_bench2_single_process.err
