#!/bin/bash
# round 2, visit d: native exchange + bucketed table: parity, bench lines, ncu of the bucket kernels, the full default bench line
mkdir -p gpurun_out
exec > gpurun_out/r2d.log 2>&1
set -x
date +%s
timeout 900 python -m pytest tests/test_gpu_exchange.py tests/test_gpu_highcard_variants.py tests/test_gpu_ingest.py -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/r2d_pytest_new.log 2>&1; echo "pytest new exit $?"
tail -25 gpurun_out/r2d_pytest_new.log
date +%s
b() {  # name, bench args...
  name=$1; shift
  timeout 500 python bench.py "$@" > gpurun_out/r2d_bench_$name.json 2> gpurun_out/r2d_bench_$name.err; echo "bench $name exit $?"
  cat gpurun_out/r2d_bench_$name.json; tail -5 gpurun_out/r2d_bench_$name.err
  date +%s
}
b highcard --workload highcard --steps 5 --no-e2e
# launch list + full captures of the bucket page kernel and the bucket emit kernel (one step each)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2d_launches_highcard.csv python bench.py --workload highcard --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/r2d_ncu_l.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pgpu_k_page -s 4 -c 1 -o gpurun_out/r2d_hc_full -f python bench.py --workload highcard --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/r2d_ncu_f.log 2>&1; echo "ncu full exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pgpu_k_emit -s 1 -c 1 -o gpurun_out/r2d_hc_emit_full -f python bench.py --workload highcard --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/r2d_ncu_e.log 2>&1; echo "ncu emit exit $?"
date +%s
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 300 -p no:cacheprovider > gpurun_out/r2d_pytest_gpu.log 2>&1; echo "pytest gpu exit $?"
tail -15 gpurun_out/r2d_pytest_gpu.log
date +%s
( time timeout 900 python bench.py > gpurun_out/r2d_bench_default.json 2> gpurun_out/r2d_bench_default.err ) 2>&1 | tail -4; echo "bench default exit $?"
cat gpurun_out/r2d_bench_default.json; tail -5 gpurun_out/r2d_bench_default.err
date +%s
