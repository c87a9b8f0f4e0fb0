#!/bin/bash
mkdir -p gpurun_out
exec > gpurun_out/r2v.log 2>&1
for inf in 2 4 3; do
  for mb in 4 3; do
    PGPU_TUNE_SWEEP_INFLIGHT=$inf PGPU_TUNE_SWEEP_MINBLOCKS=$mb timeout 400 python bench.py --workload highcard_shuffled --steps 5 --no-e2e --no-cpu > gpurun_out/sw_inf${inf}_mb${mb}.json 2>gpurun_out/sw_inf.err || tail -3 gpurun_out/sw_inf.err
    python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/sw_inf${inf}_mb${mb}.json").read().strip().splitlines()[-1])
    print("inflight $inf minblocks $mb: ms/step %.2f kernel %.2f" % (d["ms_per_step"], d["roofline"]["kernel_ms_per_step"]))
except Exception as e:
    print("inflight $inf minblocks $mb failed", e)
PY
  done
done
