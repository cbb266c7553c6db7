#!/bin/bash
# bench.py at G GPUs, N = 8192, single-process form only
G=${1:-2}; TAG=${2:-r02s}
mkdir -p gpurun_out
HYB_MULTI_MODE=single_process python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $G --steps 20 --warmup 3 > gpurun_out/${TAG}_g${G}_N8192_single_process.json 2> gpurun_out/${TAG}_g${G}.err
python - <<PY
import json
d=json.loads(open("gpurun_out/${TAG}_g${G}_N8192_single_process.json").read().strip().splitlines()[-1])
print("G=$G: ms/step", round(d["ms_per_step"],2), "e2e it/s", round(d["e2e"]["value"],2), "asm", round(d["schur_assemble_ms"],2), "reduce", round(d["schur_reduce_ms"],2), "chol", round(d["cholesky_ms"],2), {k: round(v,2) for k,v in d["phase_ms"].items()})
PY
