#!/bin/bash
mkdir -p gpurun_out
exec > gpurun_out/trip.log 2>&1
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
run() {
  w=$1; shift
  env "$@" timeout 900 python bench.py --workload $w --steps 5 --no-cpu --no-e2e 2>gpurun_out/trip_$w.err | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$w $*', d['ms_per_step'], d['roofline']['frac'], d['roofline']['avg_launch_ms'], d['gpu_launches'], d['result_rows'])"
  tail -3 gpurun_out/trip_$w.err
}
run highcard A=1
run highcard PGPU_TUNE_MINBLOCKS=2
run highcard PGPU_TUNE_MINBLOCKS=4
run highcard_shuffled A=1
run q1 A=1
run q6 A=1
