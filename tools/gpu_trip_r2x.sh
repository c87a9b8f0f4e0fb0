#!/bin/bash
# streaming runs: parity tests + launch list of one GROUP BY l_orderkey step
mkdir -p gpurun_out
exec > gpurun_out/r2x.log 2>&1
timeout 900 python -m pytest tests/test_gpu_stream.py -q -m gpu 2>&1 | tail -25
timeout 300 python bench.py --workload highcard --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/hc_x.json 2>gpurun_out/hc_x.err || tail -5 gpurun_out/hc_x.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/hc_launches.csv python bench.py --workload highcard --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/hc_ncu.log 2>&1
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/hc_launches.csv")) if len(r) > 5]
hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value"); ui = hdr.index("Metric Unit")
seq = [(r[ki][:50], float(r[vi].replace(",", "")), r[ui]) for r in rows[1:]]
for s in seq[-40:]: print(s)
PY
