#!/bin/bash
# streaming runs: parity tests, then GROUP BY l_orderkey under the rows-per-thread / occupancy knobs
mkdir -p gpurun_out
exec > gpurun_out/r2y.log 2>&1
timeout 900 python -m pytest tests/test_gpu_stream.py -q -m gpu 2>&1 | tail -8
for cfg in "1 4"; do
  set -- $cfg
  PGPU_TUNE_STREAM_SUB=$1 PGPU_TUNE_STREAM_MINBLOCKS=$2 timeout 300 python bench.py --workload highcard --steps 5 --no-e2e --no-cpu > gpurun_out/hc_y.json 2>gpurun_out/hc_y.err || tail -5 gpurun_out/hc_y.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/hc_y.json").read().strip().splitlines()[-1])
    print("sub $1 minblocks $2: ms/step %.2f kernel %.2f avg launch %.3f" % (d["ms_per_step"], d["roofline"]["kernel_ms_per_step"], d["roofline"]["avg_launch_ms"]))
except Exception as e:
    print("sub $1 minblocks $2 failed", e)
PY
done
