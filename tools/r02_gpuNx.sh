#!/bin/bash
# bench line under torchrun at N = $1 (the driver's step counts)
set -u
N=$1
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench n$N rc=$?"; tail -c 200 gpurun_out/bench_n$N.err
