#!/bin/bash
# second GPU call of round 2: ncu --set full captures of the kernels VERDICT names + launch list of the bench command
set -u
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on"
# seed filter at the N=1 and the N=8 shard size
timeout 300 python tools/prof_target.py filter 3000000000 > gpurun_out/filter_full.log 2>&1 &&
  timeout 600 $NCU -k regex:pigeon_filter -s 2 -c 1 -o gpurun_out/r02_filter_n1 python tools/prof_target.py filter 3000000000 > gpurun_out/ncu_filter1.log 2>&1
echo "filter n1 rc=$?"
timeout 300 python tools/prof_target.py filter 375000000 > gpurun_out/filter_shard.log 2>&1 &&
  timeout 600 $NCU -k regex:pigeon_filter -s 2 -c 1 -o gpurun_out/r02_filter_n8 python tools/prof_target.py filter 375000000 > gpurun_out/ncu_filter8.log 2>&1
echo "filter n8 rc=$?"
# journal materialisation
timeout 300 python tools/prof_target.py e2e 3200000000 > gpurun_out/e2e_target.log 2>&1 &&
  timeout 900 $NCU -k regex:'pack_chunks_kernel.*Journal' -s 1 -c 1 -o gpurun_out/r02_journal python tools/prof_target.py e2e 3200000000 > gpurun_out/ncu_journal.log 2>&1
echo "journal rc=$?"
# Myers kernels
timeout 300 python tools/prof_target.py myers > gpurun_out/myers_target.log 2>&1 &&
  timeout 600 $NCU -k regex:'myers_cut_kernel' -s 2 -c 1 -o gpurun_out/r02_myers_cut python tools/prof_target.py myers > gpurun_out/ncu_myers.log 2>&1
echo "myers rc=$?"
# launch list of the bench command
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_s3.json 2> gpurun_out/bench_s3.err &&
  timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02_bench_launches.csv python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
