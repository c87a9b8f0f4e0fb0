#!/bin/bash
# round 2, visit h: the allocation-free entry point — alone in a fresh process, then after the whole GPU suite
mkdir -p gpurun_out
exec > gpurun_out/r2h.log 2>&1
set -x
date +%s
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 200 -p no:cacheprovider > gpurun_out/r2h_pytest_alone.log 2>&1; echo "pytest alone exit $?"
tail -15 gpurun_out/r2h_pytest_alone.log
date +%s
timeout 1800 python -m pytest tests -m gpu -q -x --timeout 300 -p no:cacheprovider > gpurun_out/r2h_pytest_gpu.log 2>&1; echo "pytest gpu exit $?"
tail -15 gpurun_out/r2h_pytest_gpu.log
date +%s
