#!/bin/bash
# round 2, visit b: bucketed high-cardinality table — parity, the random-slot floor of HBM, bench lines, ncu of the bucket kernel
mkdir -p gpurun_out
exec > gpurun_out/r2b.log 2>&1
set -x
date +%s
timeout 900 python -m pytest tests/test_gpu_ingest.py tests/test_gpu_highcard_variants.py -m gpu -q --timeout 200 -p no:cacheprovider > gpurun_out/r2b_pytest_new.log 2>&1; echo "pytest new exit $?"
tail -30 gpurun_out/r2b_pytest_new.log
date +%s
timeout 300 tools/microbench/randtable > gpurun_out/r2b_randtable.txt 2>&1; cat gpurun_out/r2b_randtable.txt
date +%s
b() {  # name, bench args...
  name=$1; shift
  timeout 400 python bench.py "$@" > gpurun_out/r2b_bench_$name.json 2> gpurun_out/r2b_bench_$name.err; echo "bench $name exit $?"
  cat gpurun_out/r2b_bench_$name.json; tail -5 gpurun_out/r2b_bench_$name.err
  date +%s
}
b highcard --workload highcard --steps 5 --no-e2e
PGPU_TUNE_BUCKETS=0 b highcard_perkey --workload highcard --steps 5 --no-e2e --no-cpu
b highcard_shuffled --workload highcard_shuffled --steps 5 --no-e2e
PGPU_TUNE_RUNS=0 b highcard_shuffled_norun --workload highcard_shuffled --steps 5 --no-e2e --no-cpu
# launch list + full capture of the bucket page kernel (one step)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2b_launches_highcard.csv python bench.py --workload highcard --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/r2b_ncu_l.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pgpu_k_page -s 4 -c 1 -o gpurun_out/r2b_hc_full -f python bench.py --workload highcard --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/r2b_ncu_f.log 2>&1; echo "ncu full exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pgpu_k_emit -s 1 -c 1 -o gpurun_out/r2b_hc_emit_full -f python bench.py --workload highcard --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/r2b_ncu_e.log 2>&1; echo "ncu emit exit $?"
date +%s
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 200 -p no:cacheprovider > gpurun_out/r2b_pytest_gpu.log 2>&1; echo "pytest gpu exit $?"
tail -15 gpurun_out/r2b_pytest_gpu.log
date +%s
