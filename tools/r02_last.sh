#!/bin/bash
# last check of the round on one GPU: the GPU tests and the bench line of the final build
set -u
mkdir -p gpurun_out
timeout 800 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r02_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -2 gpurun_out/r02_pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_n1.json 2> gpurun_out/bench.err
echo "bench rc=$?"; tail -c 300 gpurun_out/bench.err
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
