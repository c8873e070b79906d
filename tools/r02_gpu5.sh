#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --maxfail=10 -p no:cacheprovider > gpurun_out/pytest5.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/pytest5.log
timeout 600 python tools/stream_ab.py 3200000000 67108864 33554432 16777216 > gpurun_out/stream_ab5.log 2>&1
echo "stream rc=$?"; cat gpurun_out/stream_ab5.log | tail -8
SPMB_UPLOAD_TRACE=1 timeout 300 python tools/stream_trace.py 67108864 8 > gpurun_out/stream_trace5.log 2>&1
grep -v "mover" gpurun_out/stream_trace5.log | tail -30
timeout 600 python tools/c5_shard.py --out gpurun_out/c5_shards5.json > gpurun_out/c5_shards5.log 2>&1
echo "c5 rc=$?"; tail -5 gpurun_out/c5_shards5.log | head -4
timeout 600 python tools/c5_shard.py --steps 4 --only 8 --out gpurun_out/c5_tmp.json > /dev/null 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/c5_n8_launches.csv python tools/c5_shard.py --steps 4 --only 8 --out gpurun_out/c5_tmp.json > /dev/null 2>&1
echo "c5 launches rc=$?"
