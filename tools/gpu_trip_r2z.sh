#!/bin/bash
# ncu --set full of one streaming-runs kernel ($1 = count | write | emit; one launch over the whole page). (While the path was one
# pass this script captured pgpu_k_runs_stream: the barrier / look-back findings in profiles/r02_stream.md come from those captures.)
K=${1:-write}
mkdir -p gpurun_out
exec > gpurun_out/r2z.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:^pgpu_k_runs_$K\$ --launch-skip 1 -c 1 -o gpurun_out/runs_${K}_full -f python bench.py --workload highcard --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_runs_$K.log 2>&1
ls -la gpurun_out/*.ncu-rep
