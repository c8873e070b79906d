#!/bin/bash
# host/GPU timeline of the pipelined passes of C5 at shard sizes 1/2 and 1/8 (one GPU)
set -u
mkdir -p gpurun_out
SPMB_BENCH_TRACE=1 timeout 300 python tools/c5_shard.py --only 2 --steps 10 --out gpurun_out/c5_t2.json > gpurun_out/c5_t2.log 2> gpurun_out/c5_t2.trace
echo "rc=$?"; tail -2 gpurun_out/c5_t2.log | head -1
SPMB_BENCH_TRACE=1 timeout 300 python tools/c5_shard.py --only 8 --steps 10 --out gpurun_out/c5_t8.json > gpurun_out/c5_t8.log 2> gpurun_out/c5_t8.trace
echo "rc=$?"; tail -2 gpurun_out/c5_t8.log | head -1
