#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 800 python -m pytest tests -m gpu -q -x -p no:cacheprovider 2>&1 | tail -3
SPMB_BENCH_TRACE=1 timeout 300 python tools/c5_shard.py --only 2 --steps 10 --out gpurun_out/c5_t2.json > gpurun_out/c5_t2.log 2> gpurun_out/c5_t2.trace
tail -2 gpurun_out/c5_t2.log | head -1
SPMB_BENCH_TRACE=1 timeout 300 python tools/c5_shard.py --only 8 --steps 10 --out gpurun_out/c5_t8.json > gpurun_out/c5_t8.log 2> gpurun_out/c5_t8.trace
tail -2 gpurun_out/c5_t8.log | head -1
