#!/bin/bash
# Final validation visit: the whole GPU suite, smoke(), the bench line and the reference arm.
mkdir -p gpurun_out
TAG=${1:-r02_final}
python -m pytest tests -m gpu -q -s 2>&1 | tail -120 > gpurun_out/${TAG}_tests.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${TAG}_smoke.log 2>&1
python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
tail -4 gpurun_out/${TAG}_tests.log; tail -2 gpurun_out/${TAG}_smoke.log; head -c 900 gpurun_out/${TAG}_bench.json; echo; head -c 500 gpurun_out/${TAG}_bench_ref.json
