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
NCU="ncu --set full --clock-control none --import-source on"
timeout 300 python tools/prof_target.py both > gpurun_out/t_both.log 2>&1 &&
  timeout 900 $NCU -k regex:'exact_kernel|myers_kernel|myers_cut_kernel' -s 3 -c 16 -o gpurun_out/r02_exact_myers python tools/prof_target.py both > gpurun_out/ncu_both.log 2>&1
echo "ncu exact/myers rc=$?"
timeout 300 python tools/c5_shard.py --steps 4 --only 1 --out gpurun_out/c5_tmp.json > /dev/null 2>&1 &&
  timeout 900 $NCU -k regex:'pigeon_filter|pigeon_verify' -s 8 -c 2 -o gpurun_out/r02_c5_n1 python tools/c5_shard.py --steps 4 --only 1 --out gpurun_out/c5_tmp.json > gpurun_out/ncu_c5a.log 2>&1
echo "ncu c5 n1 rc=$?"
timeout 300 python tools/c5_shard.py --steps 4 --only 8 --out gpurun_out/c5_tmp.json > /dev/null 2>&1 &&
  timeout 900 $NCU -k regex:'pigeon_filter|pigeon_verify' -s 8 -c 2 -o gpurun_out/r02_c5_n8 python tools/c5_shard.py --steps 4 --only 8 --out gpurun_out/c5_tmp.json > gpurun_out/ncu_c5b.log 2>&1
echo "ncu c5 n8 rc=$?"
timeout 300 python tools/prof_target.py e2e 3200000000 > gpurun_out/r02_e2e_target.txt 2>&1 &&
  timeout 900 $NCU -k regex:'journal_derive_kernel|pack_chunks_kernel' -s 2 -c 3 -o gpurun_out/r02_journal python tools/prof_target.py e2e 3200000000 > gpurun_out/ncu_journal.log 2>&1
echo "ncu journal rc=$?"
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_s3.json 2> gpurun_out/bench_s3.err &&
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r02_bench_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
