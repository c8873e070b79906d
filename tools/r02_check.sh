#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 800 python -m pytest tests -m gpu -q -x -p no:cacheprovider 2>&1 | tail -3
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-myers --no-small --no-c5 --no-multi --no-stream --no-pageable --e2e-steps 1 > gpurun_out/bench_c4_quick.json 2> gpurun_out/bench_c4_quick.err
echo "rc=$?"; tail -c 300 gpurun_out/bench_c4_quick.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_c4_quick.json').read().strip().splitlines()[-1])
c=d.get('c4') or {}
print(' c4 ms/step', c.get('ms_per_step'), 'kernels', c.get('kernels_ms'), 'sync', c.get('synchronous_call_ms'), c.get('synchronous_call_kernels_ms'), 'expand', c.get('expand_ms'), c.get('parity'), c.get('pipelined_hits_equal_synchronous'), c.get('one_off'))
PY
