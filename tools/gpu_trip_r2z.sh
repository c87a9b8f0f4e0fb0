#!/bin/bash
# ncu --set full of the streaming-runs kernel (one steady-state launch)
mkdir -p gpurun_out
exec > gpurun_out/r2z.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pgpu_k_runs_stream --launch-skip 12 -c 1 -o gpurun_out/stream_full -f python bench.py --workload highcard --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_stream.log 2>&1
ls -la gpurun_out/*.ncu-rep
