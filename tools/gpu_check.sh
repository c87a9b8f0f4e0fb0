#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench lines for every workload, the ncu launch list and full captures of the
# dominant kernels. Everything lands in gpurun_out/ (scratch); summaries are copied into profiles/ by tools/collect_profiles.sh.
mkdir -p gpurun_out
exec > gpurun_out/gpu_check.log 2>&1
set -x
nvidia-smi -L; nproc; free -g | head -2; which java javac mvn; ls /usr/lib/jvm
timeout 900 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 1200 python bench.py > gpurun_out/bench_q1.json 2> gpurun_out/bench_q1.err; echo "bench q1 exit $?"
cat gpurun_out/bench_q1.json
for w in q6 highcard highcard_shuffled q1dict; do
  timeout 600 python bench.py --workload $w --steps 5 --no-cpu > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w exit $?"
  cat gpurun_out/bench_$w.json
done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"
cat gpurun_out/bench_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_q1.csv \
  python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pgpu_k_page_staged -s 2 -c 1 -f -o gpurun_out/q1_full \
  python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --rows 200000000 > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pgpu_k_page -s 1 -c 1 -f -o gpurun_out/hc_full \
  python bench.py --workload highcard --steps 1 --warmup 1 --no-e2e --no-cpu --rows 200000000 > gpurun_out/ncu_full_hc.log 2>&1; echo "ncu full hc exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pgpu_k_page -s 1 -c 1 -f -o gpurun_out/q6_full \
  python bench.py --workload q6 --steps 1 --warmup 1 --no-e2e --no-cpu --rows 200000000 > gpurun_out/ncu_full_q6.log 2>&1; echo "ncu full q6 exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pgpu_k_page_staged -s 2 -c 1 -f -o gpurun_out/q1dict_full \
  python bench.py --workload q1dict --steps 1 --warmup 1 --no-e2e --no-cpu --rows 200000000 > gpurun_out/ncu_full_q1dict.log 2>&1; echo "ncu full q1dict exit $?"
timeout 300 python tools/parquet_ingest_bench.py > gpurun_out/parquet_ingest.json 2> gpurun_out/parquet_ingest.err; echo "parquet ingest exit $?"
cat gpurun_out/parquet_ingest.json; tail -3 gpurun_out/parquet_ingest.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_highcard.csv \
  python bench.py --workload highcard --steps 1 --warmup 1 --no-e2e --no-cpu > /dev/null 2>&1; echo "ncu launches highcard exit $?"
ls -la gpurun_out
