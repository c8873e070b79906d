#!/bin/bash
# evidence run of the final build on one B200: tests, both bench arms, launch list, ncu --set full of every hot kernel,
# the adversarial set, the C5 shard curve, the stream sweep
set -u
mkdir -p gpurun_out
rm -f gpurun_out/*.ncu-rep
timeout 1200 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r02_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -2 gpurun_out/r02_pytest_gpu.log
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/bench_ref.err
echo "ref arm rc=$?"
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_bench_n1.json 2> gpurun_out/bench.err
echo "bench rc=$?"; tail -c 300 gpurun_out/bench.err
timeout 600 python tools/adversarial.py --out gpurun_out/r02_adversarial.json > gpurun_out/adversarial.log 2>&1
echo "adversarial rc=$?"
timeout 600 python tools/c5_shard.py --out gpurun_out/r02_c5_shards.json > gpurun_out/c5.log 2>&1
echo "c5 rc=$?"; tail -5 gpurun_out/c5.log | head -4
timeout 600 python tools/stream_ab.py 3200000000 67108864 33554432 268435456 > gpurun_out/r02_stream_ab.txt 2>&1
echo "stream rc=$?"; tail -4 gpurun_out/r02_stream_ab.txt
