#!/bin/bash
# first GPU call of round 2: tests, one bench run, the adversarial set, the C5 shard curve
set -u
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt 2>&1
nproc >> gpurun_out/gpus.txt; free -g >> gpurun_out/gpus.txt
timeout 900 python -m pytest tests -m gpu -q --maxfail=30 -x -p no:cacheprovider > gpurun_out/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -5 gpurun_out/pytest.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?"; tail -c 600 gpurun_out/bench.err
timeout 600 python tools/adversarial.py --out gpurun_out/adversarial.json > gpurun_out/adversarial.log 2>&1
echo "adversarial rc=$?"
timeout 600 python tools/c5_shard.py --out gpurun_out/c5_shards.json > gpurun_out/c5_shards.log 2>&1
echo "c5 rc=$?"; tail -3 gpurun_out/c5_shards.log
