#!/bin/bash
# e2e leg only, N ranks: default vs write-combined staging slots, thread counts
N=${1:-1}
mkdir -p gpurun_out
exec > gpurun_out/r2_e2e$N.log 2>&1
run() {
  name=$1; shift
  if [ $N -gt 1 ]; then L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514"; else L=python; fi
  env "$@" timeout 600 $L bench.py --gpus $N --workload q1 --steps 3 --warmup 3 --no-cpu --e2e-rows 256000000 > gpurun_out/r2_e2e${N}_$name.json 2>/dev/null
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2_e2e${N}_$name.json").read().strip().splitlines()[-1]); e=d["e2e"]
    print("$name N=$N e2e %.3f G rows/s  %.1f GB/s  h2d ceiling %.1f  host copy ceiling %.1f  threads %s" % (e["value"]/1e9, e["gb_per_s"], e["h2d_ceiling_gbs"], e["host_copy_ceiling_gbs"], e["config"]["copy_threads"]))
except Exception as ex:
    print("$name failed", ex)
PY
}
nproc
run default A=1
run wc PGPU_STAGE_WC=1
run wc_t4 PGPU_STAGE_WC=1 PGPU_COPY_THREADS=4
run t4 PGPU_COPY_THREADS=4
run stage16 PGPU_STAGE_MB=16
