#!/bin/bash
# round 2, visit j: partition sweep — parity tests, then GROUP BY l_orderkey (clustered / shuffled) with and without the sweep
mkdir -p gpurun_out
exec > gpurun_out/r2j.log 2>&1
set -x
date +%s
timeout 900 python -m pytest tests/test_gpu_sweep.py tests/test_merge_pages.py -m gpu -x -q 2>&1 | tail -15
for wl in highcard highcard_shuffled; do
  timeout 400 python bench.py --workload $wl --steps 5 --no-e2e --no-cpu > gpurun_out/sweep_${wl}_on.json 2> gpurun_out/sweep_${wl}_on.err; echo "$wl on exit $?"; tail -2 gpurun_out/sweep_${wl}_on.err
  PGPU_TUNE_SWEEP=0 timeout 400 python bench.py --workload $wl --steps 5 --no-e2e --no-cpu > gpurun_out/sweep_${wl}_off.json 2> gpurun_out/sweep_${wl}_off.err; echo "$wl off exit $?"
done
python - <<'PY'
import json
for wl in ("highcard", "highcard_shuffled"):
    for m in ("on", "off"):
        try:
            d = json.loads(open("gpurun_out/sweep_%s_%s.json" % (wl, m)).read().strip().splitlines()[-1])
            print(wl, m, "ms/step %.2f kernel %.2f verified %s launches %s" % (d["ms_per_step"], d["roofline"]["kernel_ms_per_step"], d.get("verified"), d.get("gpu_launches")))
        except Exception as e:
            print(wl, m, "failed", e)
PY
date +%s
