#!/bin/bash
# ncu evidence of the final build on one B200: --set full of every hot kernel (reports are read ON THE BOX into
# gpurun_out/r02_ncu_full_raw.csv and deleted: gpurun brings back 64 MiB at most), then the bench command's launch list
set -u
mkdir -p gpurun_out
rm -f gpurun_out/*.ncu-rep
NCU="ncu --set full --clock-control none --import-source on"
timeout 300 python tools/prof_target.py both > gpurun_out/t_both.log 2>&1 &&
  timeout 900 $NCU -k regex:'exact_kernel|myers_kernel|myers_cut_kernel' -s 3 -c 12 -o gpurun_out/r02_exact_myers python tools/prof_target.py both > gpurun_out/ncu_both.log 2>&1
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
python tools/ncu_extract.py gpurun_out/*.ncu-rep > gpurun_out/r02_ncu_extract.log 2>&1; cp profiles/r02_ncu_full_raw.csv gpurun_out/; rm -f gpurun_out/*.ncu-rep
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_s3.json 2> gpurun_out/bench_s3.err &&
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r02_bench_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
# the later legs of the bench command (Myers C3, JST C4, C5, C1/C2) -- the first list ends inside the stream leg
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu --no-stream --no-multi --e2e-steps 1 > gpurun_out/bench_s3b.json 2> gpurun_out/bench_s3b.err &&
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -s 2500 -c 6000 --csv --log-file gpurun_out/r02_bench_launches_late.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-stream --no-multi --e2e-steps 1 > gpurun_out/ncu_bench_b.log 2>&1
echo "late launch list rc=$?"

