#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --maxfail=10 -p no:cacheprovider > gpurun_out/pytest7.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/pytest7.log
timeout 600 python tools/stream_ab.py 3200000000 67108864 33554432 16777216 > gpurun_out/stream_ab7.log 2>&1
echo "stream rc=$?"; cat gpurun_out/stream_ab7.log | tail -8
timeout 600 python tools/c5_shard.py --out gpurun_out/c5_shards7.json > gpurun_out/c5_shards7.log 2>&1
echo "c5 rc=$?"; tail -5 gpurun_out/c5_shards7.log | head -4
