#!/bin/bash
# visit 4: the whole GPU suite (with the wire-format ingest and the TPC-H golden-answer tests), then the bench on dbgen rows
mkdir -p gpurun_out
exec > gpurun_out/trip4.log 2>&1
set -x
timeout 400 python -m pytest tests -m gpu -x -q --timeout 200 > gpurun_out/pytest_gpu_r01f.log 2>&1; echo "pytest exit $?"
tail -15 gpurun_out/pytest_gpu_r01f.log
for w in q1 q6 highcard; do
  timeout 120 python bench.py --data dbgen --sf 100 --workload $w --steps 5 --no-cpu --no-e2e > gpurun_out/bench_dbgen_$w.json 2> gpurun_out/bench_dbgen_$w.err; echo "bench dbgen $w exit $?"
  cat gpurun_out/bench_dbgen_$w.json; tail -3 gpurun_out/bench_dbgen_$w.err
done
