#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/fuse.log
: > $O
python -m pytest tests -m gpu -q 2>&1 | tail -15 >> $O
cat $O; head -c 1500 gpurun_out/fuse_bench.json
