#!/bin/bash
mkdir -p gpurun_out
exec > gpurun_out/r2q.log 2>&1
for wt in 1024 512 256 2048; do
  for k in 0 128 512; do
    if [ $k = 0 ]; then unset PGPU_TUNE_SWEEP_K; else export PGPU_TUNE_SWEEP_K=$k; fi
    PGPU_TUNE_SWEEP_WT=$wt timeout 400 python bench.py --workload highcard_shuffled --steps 5 --no-e2e --no-cpu > gpurun_out/sw_wt${wt}_k${k}.json 2>/dev/null
    python - <<PY
import json
d = json.loads(open("gpurun_out/sw_wt${wt}_k${k}.json").read().strip().splitlines()[-1])
print("WT $wt K $k: ms/step %.2f kernel %.2f" % (d["ms_per_step"], d["roofline"]["kernel_ms_per_step"]))
PY
  done
done
