#!/bin/bash
# round 2, visit l: partition sweep, scatter tile 4096, clustering decision; full ncu of the sweep kernel
mkdir -p gpurun_out
exec > gpurun_out/r2l.log 2>&1
set -x
timeout 900 python -m pytest tests/test_gpu_sweep.py -m gpu -x -q 2>&1 | tail -5
for wl in highcard highcard_shuffled; do
  timeout 400 python bench.py --workload $wl --steps 5 --no-e2e --no-cpu > gpurun_out/sweep_${wl}_on.json 2> gpurun_out/sweep_${wl}_on.err; echo "$wl on exit $?"; tail -2 gpurun_out/sweep_${wl}_on.err
done
python - <<'PY'
import json
for wl in ("highcard", "highcard_shuffled"):
    for m in ("on",):
        try:
            d = json.loads(open("gpurun_out/sweep_%s_%s.json" % (wl, m)).read().strip().splitlines()[-1])
            print(wl, m, "ms/step %.2f kernel %.2f verified %s launches %s" % (d["ms_per_step"], d["roofline"]["kernel_ms_per_step"], d.get("verified"), d.get("gpu_launches")))
        except Exception as e:
            print(wl, m, "failed", e)
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 16 --csv --log-file gpurun_out/launches_sweep_shuf3.csv python bench.py --workload highcard_shuffled --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_l_sw.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pgpu_k_page_sweep -s 10 -c 1 -f -o gpurun_out/sweep_shuf_full2 python bench.py --workload highcard_shuffled --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_swf.log 2>&1; echo exit $?
