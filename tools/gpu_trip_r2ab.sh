#!/bin/bash
# streaming runs at the sf1000 / 4 GPUs shard size (1.5 G rows on one GPU): the path does not depend on the page length
mkdir -p gpurun_out
exec > gpurun_out/r2ab.log 2>&1
timeout 200 python bench.py --workload highcard --rows 1499997427 --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench_highcard_1500m.json 2> gpurun_out/bench_highcard_1500m.err || tail -5 gpurun_out/bench_highcard_1500m.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_highcard_1500m.json").read().strip().splitlines()[-1])
print("ms/step %.2f rows/s %.4g kernel %.2f verified %s streamed %s result rows %s" % (d["ms_per_step"], d["value"], d["roofline"]["kernel_ms_per_step"], d.get("verified"), d.get("streamed_runs_per_step"), d.get("result_rows")))
PY
