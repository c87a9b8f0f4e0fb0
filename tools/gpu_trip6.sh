#!/bin/bash
# visit 6: the whole GPU suite (with the high-cardinality variant tests), then GROUP BY l_orderkey at sf100 on the variants
mkdir -p gpurun_out
exec > gpurun_out/trip6.log 2>&1
set -x
date +%s
timeout 200 python -m pytest tests -m gpu -q --timeout 100 -p no:cacheprovider > gpurun_out/pytest_gpu_r01g.log 2>&1; echo "pytest exit $?"
tail -25 gpurun_out/pytest_gpu_r01g.log
date +%s
b() {  # name, workload, data args..., env in the caller
  name=$1; shift
  timeout 70 python bench.py "$@" --steps 5 --no-cpu --no-e2e > gpurun_out/bench_hcv_$name.json 2> gpurun_out/bench_hcv_$name.err; echo "bench $name exit $?"
  cat gpurun_out/bench_hcv_$name.json; tail -3 gpurun_out/bench_hcv_$name.err
  date +%s
}
PGPU_TUNE_RUNS=1 PGPU_TUNE_LOCB=0 b runs1_loc0 --data dbgen --sf 100 --workload highcard
PGPU_TUNE_RUNS=1 PGPU_TUNE_LOCB=4 b runs1_loc4 --data dbgen --sf 100 --workload highcard
PGPU_TUNE_RUNS=1 PGPU_TUNE_LOCB=0 b shuf_runs1_loc0 --workload highcard_shuffled
PGPU_TUNE_RUNS=0 PGPU_TUNE_LOCB=4 b runs0_loc4 --data dbgen --sf 100 --workload highcard
PGPU_TUNE_RUNS=1 PGPU_TUNE_LOCB=4 b shuf_runs1_loc4 --workload highcard_shuffled
