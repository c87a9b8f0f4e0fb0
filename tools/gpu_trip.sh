#!/bin/bash
mkdir -p gpurun_out
exec > gpurun_out/trip.log 2>&1
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
for w in q1 highcard; do
  timeout 900 python bench.py --workload $w --steps 5 --no-cpu 2>gpurun_out/trip_$w.err > gpurun_out/trip_$w.json; tail -1 gpurun_out/trip_$w.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$w', d['ms_per_step'], d['roofline']['frac'], d['gpu_launches'], d['e2e'])"
  tail -3 gpurun_out/trip_$w.err
done
