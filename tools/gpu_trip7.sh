#!/bin/bash
# visit 7 (short: whatever box time is left): bench on dbgen rows, the high-cardinality variants, then the newest GPU tests
mkdir -p gpurun_out
exec > gpurun_out/trip7.log 2>&1
set -x
date +%s
b() {  # name, bench args... (tuning knobs from the caller's environment)
  name=$1; shift
  timeout 70 python bench.py "$@" --steps 5 --no-cpu --no-e2e > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "bench $name exit $?"
  cat gpurun_out/bench_$name.json; tail -3 gpurun_out/bench_$name.err
  date +%s
}
b dbgen_q1 --data dbgen --sf 100 --workload q1
b dbgen_highcard --data dbgen --sf 100 --workload highcard
PGPU_TUNE_RUNS=1 PGPU_TUNE_LOCB=4 b hcv_runs1_loc4 --data dbgen --sf 100 --workload highcard
PGPU_TUNE_RUNS=0 PGPU_TUNE_LOCB=4 b hcv_runs0_loc4 --data dbgen --sf 100 --workload highcard
PGPU_TUNE_RUNS=1 PGPU_TUNE_LOCB=0 b hcv_runs1_loc0 --data dbgen --sf 100 --workload highcard
b dbgen_q6 --data dbgen --sf 100 --workload q6
timeout 300 python -m pytest tests -m gpu -q --timeout 100 -p no:cacheprovider > gpurun_out/pytest_gpu_r01g.log 2>&1; echo "pytest exit $?"
tail -25 gpurun_out/pytest_gpu_r01g.log
date +%s
