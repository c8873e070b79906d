#!/bin/bash
# 8 ranks, headline + C4 + C5 only, the driver's step counts
set -u
mkdir -p gpurun_out
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu --no-myers --no-small --no-c5 --no-multi --no-stream --no-pageable --e2e-steps 1 > gpurun_out/bench_n8_quick.json 2> gpurun_out/bench_n8_quick.err
echo "rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_n8_quick.json').read().strip().splitlines()[-1])
print(' headline Gbp/s', round(d['value'],1), 'ms/step', round(d['ms_per_step'],4), 'kernel', round(d['roofline']['kernel_ms'],4), 'gather_ms', d['detail'].get('final_gather_ms_this_rank'))
c=d.get('c5') or {}
print(' c5 ms/step', c.get('ms_per_step'), 'kernels', c.get('kernels_ms'), 'gather_ms', c.get('final_gather_ms_this_rank'), 'hits', c.get('hits'), c.get('merged_hits_per_pass_rank0'))
c=d.get('c4') or {}
print(' c4 ms/step', c.get('ms_per_step'), 'kernels', c.get('kernels_ms'), c.get('parity'), json.dumps(c.get('per_rank_ms')))
PY
