#!/bin/bash
# Sweep of the Q1 kernel tuning knobs (codegen.cpp / pipeline.cpp getenv hooks) on the GPU box; prints ms/step per combination.
mkdir -p gpurun_out
exec > gpurun_out/tune_q1.log 2>&1
W=${1:-q1}
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
run() {
  r=$(env "$@" timeout 300 python bench.py --workload $W --steps 5 --warmup 3 --no-e2e --no-cpu 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['frac'])" 2>&1 | tail -1)
  echo "W=$W $* -> $r"
}
run PGPU_TUNE_RG=4
run PGPU_TUNE_RG=3
run PGPU_TUNE_RG=2
run PGPU_TUNE_RG=0
run PGPU_TUNE_RG=4 PGPU_TUNE_STAGES=2
run PGPU_TUNE_RG=4 PGPU_TUNE_STAGES=3
run PGPU_TUNE_NOSTAGE=1 PGPU_TUNE_RG=0
