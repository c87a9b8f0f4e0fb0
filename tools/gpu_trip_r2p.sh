#!/bin/bash
# round 2, visit p: the evidence run after the partition sweep — default bench line (all configs, verified), reference arm, launch lists,
# one full ncu capture per workload's dominant kernel at the benchmarked size, the microbenchmarks, the ingest sweeps, Parquet ingest.
mkdir -p gpurun_out
exec > gpurun_out/r2p.log 2>&1
set -x
date +%s
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
( time timeout 1200 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err ) 2>&1 | tail -4; echo "bench default exit $?"
tail -3 gpurun_out/bench_default.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"
date +%s
cap() {  # name kernel-regex skip workload [env...]
  name=$1; kern=$2; skip=$3; wl=$4; shift 4
  env "$@" timeout 900 ncu --set full --clock-control none --import-source on -k regex:$kern -s $skip -c 1 -f -o gpurun_out/${name}_full \
    python bench.py --workload $wl --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_${name}.log 2>&1; echo "ncu $name exit $?"
}
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_q1.csv python bench.py --workload q1 --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_l_q1.log 2>&1; echo "ncu launches q1 exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_highcard.csv python bench.py --workload highcard --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_l_hc.log 2>&1; echo "ncu launches highcard exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct --clock-control none -c 60 --csv --log-file gpurun_out/launches_highcard_shuffled.csv python bench.py --workload highcard_shuffled --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_l_hcs.log 2>&1; echo "ncu launches highcard_shuffled exit $?"
cap q1 '^pgpu_k_page_staged$' 12 q1 A=1
cap q6 '^pgpu_k_page$' 2 q6 A=1
cap q1dict '^pgpu_k_page_staged$' 12 q1dict A=1
cap highcard '^pgpu_k_page$' 22 highcard A=1
cap highcard_shuffled '^pgpu_k_page_sweep$' 12 highcard_shuffled A=1
cap highcard_shuffled_scatter '^pgpu_k_part_scatter$' 1 highcard_shuffled A=1
cap highcard_shuffled_plain '^pgpu_k_page$' 12 highcard_shuffled PGPU_TUNE_SWEEP=0
cap highcard_emit '^pgpu_k_emit$' 1 highcard A=1
date +%s
PGPU_TUNE_SWEEP=0 timeout 300 python bench.py --workload highcard_shuffled --steps 5 --no-e2e --no-cpu > gpurun_out/bench_highcard_shuffled_nosweep.json 2>/dev/null
PGPU_TUNE_BUCKETS=1 PGPU_TUNE_BCAS=0 timeout 300 python bench.py --workload highcard --steps 5 --no-e2e --no-cpu > gpurun_out/bench_highcard_buckets.json 2>/dev/null
timeout 300 tools/microbench/randtable > gpurun_out/randtable.txt 2>&1
PAGES=32 timeout 600 python tools/parquet_ingest_bench.py > gpurun_out/parquet_ingest.json 2> gpurun_out/parquet_ingest.err; echo "parquet exit $?"; tail -2 gpurun_out/parquet_ingest.err
for t in 16 8 4 2 1; do PGPU_COPY_THREADS=$t timeout 300 python tools/ingest_bench.py --rows 128000000 --page-rows 1000000,65536,4096 --steps 3 > gpurun_out/ingest_t$t.json 2>/dev/null; done
PGPU_COPY_THREADS=16 timeout 300 python tools/ingest_bench.py --rows 128000000 --page-rows 1000000 --steps 3 --serialized > gpurun_out/ingest_serialized.json 2>/dev/null
date +%s
ls -la gpurun_out | head -80
