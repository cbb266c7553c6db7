#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/fuse.log
: > $O
python -m pytest tests -m gpu -q 2>&1 | tail -15 >> $O
python bench.py --steps 20 --warmup 3 > gpurun_out/fuse_bench.json 2>> $O
HYB_SPEC_SEARCH=0 python bench.py --steps 20 --warmup 3 > gpurun_out/fuse_bench_seq.json 2>> $O
python tools/run_cfg.py cfg1 >> $O 2>&1
HYB_SPEC_SEARCH=0 python tools/run_cfg.py cfg1 >> $O 2>&1
python tools/run_cfg.py cfg4s >> $O 2>&1
HYB_SPEC_SEARCH=0 python tools/run_cfg.py cfg4s >> $O 2>&1
cat $O
