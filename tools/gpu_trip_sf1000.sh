#!/bin/bash
# sf1000 on N GPUs of one box: 5 999 989 709 lineitem rows split over the ranks (N = 8: 749 998 714 rows per GPU, N = 4: 1 499 997 428)
N=${1:-8}
ROWS=$(( 5999989709 / N ))
mkdir -p gpurun_out
exec > gpurun_out/r2_sf1000_n$N.log 2>&1
set -x
date +%s
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 --rows $ROWS --no-e2e > gpurun_out/r2_sf1000_n${N}.json 2> gpurun_out/r2_sf1000_n${N}.err; echo "exit $?"
tail -5 gpurun_out/r2_sf1000_n${N}.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2_sf1000_n${N}.json").read().strip().splitlines()[-1])
    print("sf1000 q1 N=$N rows/gpu $ROWS ms/step", round(d["ms_per_step"],3), "value", d["value"])
    for k,v in d.get("configs",{}).items():
        print(k, "ms/step", round(v["ms_per_step"],3) if "ms_per_step" in v else v, v.get("value"))
except Exception as e:
    print("parse failed", e)
PY
date +%s
