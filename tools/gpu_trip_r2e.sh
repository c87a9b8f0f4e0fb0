#!/bin/bash
# round 2, visit e: bucketed table with probe-chosen bucket offset + CAS-first leaders: parity, A/B bench lines, steady-state ncu
mkdir -p gpurun_out
exec > gpurun_out/r2e.log 2>&1
set -x
date +%s
timeout 900 python -m pytest tests/test_gpu_highcard_variants.py tests/test_gpu_exchange.py tests/test_abi.py -q --timeout 300 -p no:cacheprovider > gpurun_out/r2e_pytest_new.log 2>&1; echo "pytest new exit $?"
tail -12 gpurun_out/r2e_pytest_new.log
date +%s
b() {  # name, bench args...
  name=$1; shift
  timeout 500 python bench.py "$@" > gpurun_out/r2e_bench_$name.json 2> gpurun_out/r2e_bench_$name.err; echo "bench $name exit $?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2e_bench_$name.json").read().strip().splitlines()[-1])
    print("$name", "ms/step", round(d["ms_per_step"],3), "kernel ms", round(d["roofline"]["kernel_ms_per_step"],3), "launches/step", d["roofline"]["launches_per_step"], "cap", d["table_capacity_slots"], "rows", d["result_rows"], "verified", d.get("verified"))
except Exception as e:
    print("$name failed", e)
PY
  tail -3 gpurun_out/r2e_bench_$name.err
}
b highcard --workload highcard --steps 5 --no-e2e
PGPU_TUNE_BCAS=0 b highcard_nocas --workload highcard --steps 5 --no-e2e --no-cpu
PGPU_TUNE_MINBLOCKS=4 b highcard_mb4 --workload highcard --steps 5 --no-e2e --no-cpu
PGPU_TUNE_MINBLOCKS=2 b highcard_mb2 --workload highcard --steps 5 --no-e2e --no-cpu
b highcard_e2e --workload highcard --steps 3 --no-cpu --e2e-rows 128000000
date +%s
# steady-state captures (third step): page kernel and emit kernel of the bucketed table
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/r2e_launches_highcard.csv python bench.py --workload highcard --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/r2e_ncu_l.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pgpu_k_page -s 22 -c 1 -o gpurun_out/r2e_hc_full -f python bench.py --workload highcard --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/r2e_ncu_f.log 2>&1; echo "ncu full exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pgpu_k_emit -s 2 -c 1 -o gpurun_out/r2e_hc_emit_full -f python bench.py --workload highcard --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/r2e_ncu_e.log 2>&1; echo "ncu emit exit $?"
date +%s
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 300 -p no:cacheprovider > gpurun_out/r2e_pytest_gpu.log 2>&1; echo "pytest gpu exit $?"
tail -8 gpurun_out/r2e_pytest_gpu.log
date +%s
