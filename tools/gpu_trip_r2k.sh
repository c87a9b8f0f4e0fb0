#!/bin/bash
# round 2, visit k: partition sweep with warp tickets; run-compaction vs direct update loop
mkdir -p gpurun_out
exec > gpurun_out/r2k.log 2>&1
set -x
timeout 900 python -m pytest tests/test_gpu_sweep.py -m gpu -x -q 2>&1 | tail -5
PGPU_TUNE_SWEEP_DIRECT=1 timeout 900 python -m pytest tests/test_gpu_sweep.py -m gpu -x -q 2>&1 | tail -5
for wl in highcard highcard_shuffled; do
  timeout 400 python bench.py --workload $wl --steps 5 --no-e2e --no-cpu > gpurun_out/sweep_${wl}_on.json 2> gpurun_out/sweep_${wl}_on.err; echo "$wl on exit $?"; tail -2 gpurun_out/sweep_${wl}_on.err
  PGPU_TUNE_SWEEP_DIRECT=1 timeout 400 python bench.py --workload $wl --steps 5 --no-e2e --no-cpu > gpurun_out/sweep_${wl}_direct.json 2> gpurun_out/sweep_${wl}_direct.err; echo "$wl direct exit $?"
done
python - <<'PY'
import json
for wl in ("highcard", "highcard_shuffled"):
    for m in ("on", "direct"):
        try:
            d = json.loads(open("gpurun_out/sweep_%s_%s.json" % (wl, m)).read().strip().splitlines()[-1])
            print(wl, m, "ms/step %.2f kernel %.2f verified %s launches %s" % (d["ms_per_step"], d["roofline"]["kernel_ms_per_step"], d.get("verified"), d.get("gpu_launches")))
        except Exception as e:
            print(wl, m, "failed", e)
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 30 --csv --log-file gpurun_out/launches_sweep_shuf2.csv python bench.py --workload highcard_shuffled --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_l_sw.log 2>&1
PGPU_TUNE_SWEEP_DIRECT=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 30 --csv --log-file gpurun_out/launches_sweep_shuf2d.csv python bench.py --workload highcard_shuffled --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_l_sw.log 2>&1
