#!/bin/bash
# 8-GPU box: the single-process multi-GPU driver at 1/2/4/8 devices, then the bench under torchrun at N=8 (the driver's step counts)
set -u
mkdir -p gpurun_out
nvidia-smi -L | wc -l; nproc; free -g | head -2
timeout 120 python -m pytest tests/test_gpu_multi.py -m gpu -q -p no:cacheprovider 2>&1 | tail -2
timeout 200 python tools/multi_ab.py --out gpurun_out/multi_ab_n8.json 2>&1 | tail -5
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err
echo "bench n8 rc=$?"; tail -c 300 gpurun_out/bench_n8.err
