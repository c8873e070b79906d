#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --maxfail=10 -p no:cacheprovider > gpurun_out/pytest4.log 2>&1
echo "pytest rc=$?"; tail -8 gpurun_out/pytest4.log
timeout 600 python tools/stream_ab.py 3200000000 67108864 33554432 268435456 > gpurun_out/stream_ab4.log 2>&1
echo "stream rc=$?"; cat gpurun_out/stream_ab4.log | tail -8
timeout 600 python tools/c5_shard.py --out gpurun_out/c5_shards4.json > gpurun_out/c5_shards4.log 2>&1
echo "c5 rc=$?"; tail -5 gpurun_out/c5_shards4.log | head -4
timeout 300 python tools/prof_target.py e2e 3200000000 > gpurun_out/e2e_target4.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'journal_derive_kernel|pack_chunks_kernel' -s 2 -c 3 -o gpurun_out/r02_journal python tools/prof_target.py e2e 3200000000 > gpurun_out/ncu_journal4.log 2>&1
echo "journal rc=$?"; cat gpurun_out/e2e_target4.log
