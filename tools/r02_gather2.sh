#!/bin/bash
# where the fixed cost of the N>1 step goes: headline + C5 only, 2 ranks, two step counts; shm gather vs NCCL gather
set -u
mkdir -p gpurun_out
for G in shm nccl; do
for K in 10 20; do
SPMB_BENCH_GATHER=$G timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps $K --warmup 3 --no-cpu --no-myers --no-small --no-c4 --no-multi --no-stream --no-pageable --e2e-steps 1 > gpurun_out/bench_n2_${G}_k$K.json 2> gpurun_out/bench_n2_${G}_k$K.err
echo "gather=$G K=$K rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_n2_${G}_k$K.json').read().strip().splitlines()[-1])
print(' headline Gbp/s', round(d['value'],1), 'ms/step', round(d['ms_per_step'],4), 'kernel', round(d['roofline']['kernel_ms'],4), 'gather_ms', d['detail'].get('final_gather_ms_this_rank'))
c=d.get('c5') or {}
print(' c5 ms/step', c.get('ms_per_step'), 'kernels', c.get('kernels_ms'), 'gather_ms', c.get('final_gather_ms_this_rank'), 'hits', c.get('hits'), c.get('merged_hits_per_pass_rank0'))
PY
done; done 2>&1 | tee gpurun_out/r02_gather_n2.txt
