#!/bin/bash
mkdir -p gpurun_out
exec > gpurun_out/r2m.log 2>&1
for mb in 3 4 5 6; do
  for d in 0 1; do
    PGPU_TUNE_SWEEP_MINBLOCKS=$mb PGPU_TUNE_SWEEP_DIRECT=$d timeout 400 python bench.py --workload highcard_shuffled --steps 5 --no-e2e --no-cpu > gpurun_out/sw_mb${mb}_d${d}.json 2>/dev/null
    python - <<PY
import json
d = json.loads(open("gpurun_out/sw_mb${mb}_d${d}.json").read().strip().splitlines()[-1])
print("minblocks $mb direct $d: ms/step %.2f kernel %.2f" % (d["ms_per_step"], d["roofline"]["kernel_ms_per_step"]))
PY
  done
done
