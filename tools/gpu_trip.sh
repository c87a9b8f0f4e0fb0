#!/bin/bash
mkdir -p gpurun_out
exec > gpurun_out/trip.log 2>&1
PGPU_TUNE_SUB=2 timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
PGPU_TUNE_SUB=2 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
run() {
  w=$1; shift
  env "$@" timeout 900 python bench.py --workload $w --steps 10 --no-cpu --no-e2e 2>gpurun_out/trip_$w.err | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$w $*', d['ms_per_step'], d['roofline']['frac'], d['roofline']['avg_launch_ms'], d['gpu_launches'], d['result_rows'])"
  tail -3 gpurun_out/trip_$w.err
}
run q1dict PGPU_TUNE_SUB=3
run q1 PGPU_TUNE_SUB=2 PGPU_TUNE_RG=3
run q6 A=1
PGPU_TUNE_SUB=2 timeout 900 ncu --set full --clock-control none --import-source on -k regex:pgpu_k_page_staged -s 2 -c 1 -f -o gpurun_out/q1s_d \
    python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --rows 200000000 2>&1 | tail -1 | cut -c1-100
