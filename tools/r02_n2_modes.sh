#!/bin/bash
set -u
mkdir -p gpurun_out
for MODE in normal noroom; do
if [ $MODE = noroom ]; then export SPMB_BENCH_SHM_NO_ROOM=1; fi
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu --no-myers --no-small --no-c4 --no-multi --no-stream --no-pageable --e2e-steps 1 > gpurun_out/bench_n2_$MODE.json 2> gpurun_out/bench_n2_$MODE.err
echo "$MODE rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_n2_$MODE.json').read().strip().splitlines()[-1])
print(' headline', round(d['value'],1), round(d['ms_per_step'],4), d['detail']['hit_lists_gathered'][:60], d['detail']['merged_hits_per_pass_rank0'])
c=d['c5']; print(' c5', c['ms_per_step'], c['kernels_ms'], c['hit_lists_gathered'][:60], c['merged_hits_per_pass_rank0'])
PY
done
