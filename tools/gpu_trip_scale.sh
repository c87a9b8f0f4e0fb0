#!/bin/bash
# multi-GPU visit: bench.py under torchrun at N = $1 (default line: Q1 headline + configs.highcard), native exchange
N=${1:-2}
mkdir -p gpurun_out
exec > gpurun_out/r2_scale$N.log 2>&1
set -x
date +%s
nvidia-smi topo -m | head -12
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2_scale${N}_default.json 2> gpurun_out/r2_scale${N}_default.err; echo "default exit $?"
tail -5 gpurun_out/r2_scale${N}_default.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2_scale${N}_default.json").read().strip().splitlines()[-1])
    print("q1 N=$N ms/step", round(d["ms_per_step"],3), "value", d["value"], "e2e", d["e2e"]["value"] if d.get("e2e") else None, "frac ceiling", d["e2e"].get("frac_of_h2d_ceiling") if d.get("e2e") else None)
    for k,v in d.get("configs",{}).items():
        print(k, "ms/step", round(v["ms_per_step"],3) if "ms_per_step" in v else v)
except Exception as e:
    print("parse failed", e)
PY
date +%s
