#!/bin/bash
# streaming runs: parity tests, then GROUP BY l_orderkey with and without the path
mkdir -p gpurun_out
exec > gpurun_out/r2w.log 2>&1
timeout 900 python -m pytest tests/test_gpu_stream.py tests/test_gpu_sweep.py -q -m gpu -x 2>&1 | tail -15
for s in 1 0; do
  PGPU_TUNE_STREAM=$s timeout 400 python bench.py --workload highcard --steps 5 --no-e2e --no-cpu > gpurun_out/hc_stream$s.json 2>gpurun_out/hc_stream.err || tail -5 gpurun_out/hc_stream.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/hc_stream$s.json").read().strip().splitlines()[-1])
    print("stream $s: ms/step %.2f kernel %.2f" % (d["ms_per_step"], d["roofline"]["kernel_ms_per_step"]), d["roofline"], d.get("config"))
except Exception as e:
    print("stream $s failed", e)
PY
done
timeout 400 python bench.py --workload highcard_shuffled --steps 5 --no-e2e --no-cpu > gpurun_out/hcs_stream1.json 2>gpurun_out/hc_stream.err || tail -5 gpurun_out/hc_stream.err
tail -c 600 gpurun_out/hcs_stream1.json
