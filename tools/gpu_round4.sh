#!/bin/bash
mkdir -p gpurun_out
TAG=${1:-r2h}
python -m pytest tests -m gpu -x -q 2>&1 | tail -40 > gpurun_out/${TAG}_tests.log
tail -6 gpurun_out/${TAG}_tests.log
for late in 0 2048 3072 4096 6144; do echo "late=$late"; HYB_POTRF_LATE=$late python tools/prof_potrf.py 8192 | tail -1; done
HYB_POTRF_LATE=3072 python tools/prof_potrf.py 4096 | tail -1; HYB_POTRF_LATE=0 python tools/prof_potrf.py 4096 | tail -1
python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; tail -2 gpurun_out/${TAG}_bench.err
HYB_HOST_STEPPER=1 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_hoststepper.json 2> gpurun_out/${TAG}_bench_hoststepper.err
python - <<PY
import json
for f in ("${TAG}_bench", "${TAG}_bench_hoststepper"):
    try:
        d=json.load(open("gpurun_out/%s.json" % f))
        print(f, "ms/step", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["value"],2), "value", round(d["value"],2), "asm", round(d["schur_assemble_ms"],2), "chol", round(d["cholesky_ms"],2), d["phase_ms"], "h2d", d["e2e"]["h2d_bytes_per_step"], d["config"]["solves"], "cpu", d.get("cpu_baseline",{}).get("value"))
    except Exception as e:
        print(f, "failed", e)
PY
