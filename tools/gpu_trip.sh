#!/bin/bash
mkdir -p gpurun_out
exec > gpurun_out/trip.log 2>&1
timeout 600 python -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout 180 > gpurun_out/trip_pytest.log 2>&1; tail -3 gpurun_out/trip_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
run() {
  w=$1; shift
  env "$@" timeout 150 python bench.py --workload $w --steps 10 --no-cpu 2>gpurun_out/trip_$w.err | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$w $*', d['ms_per_step'], d['roofline']['frac'], d['roofline']['avg_launch_ms'], d['gpu_launches'], d['result_rows'], d['e2e'])"
  tail -3 gpurun_out/trip_$w.err
}
run q1 A=1
run q6 A=1
