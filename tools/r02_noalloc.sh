#!/bin/bash
# the allocation-free entry point: its test alone in a fresh process, then behind the whole GPU suite (collected last)
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x -p no:cacheprovider > gpurun_out/r02_noalloc_alone.log 2>&1
echo "alone rc=$?"; tail -15 gpurun_out/r02_noalloc_alone.log
timeout 1200 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/r02_pytest_gpu.log 2>&1
echo "suite rc=$?"; tail -4 gpurun_out/r02_pytest_gpu.log
