#!/bin/bash
# round 2, visit aa: evidence after streaming runs — GPU suite, default bench line (all configs, verified), reference arm, launch lists with
# DRAM bytes for the two multi-kernel GROUP BY workloads, one full ncu capture of each streaming-runs kernel and of the table path beside it.
mkdir -p gpurun_out
exec > gpurun_out/r2aa.log 2>&1
set -x
date +%s
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
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
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct --clock-control none -c 60 --csv --log-file gpurun_out/launches_highcard.csv python bench.py --workload highcard --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/ncu_l_hc.log 2>&1; echo "ncu launches highcard exit $?"
cap highcard_runs_count '^pgpu_k_runs_count$' 1 highcard A=1
cap highcard_runs_write '^pgpu_k_runs_write$' 1 highcard A=1
cap highcard_runs_emit '^pgpu_k_runs_emit$' 1 highcard A=1
cap highcard_table '^pgpu_k_page$' 12 highcard PGPU_TUNE_STREAM=0
PGPU_TUNE_STREAM=0 timeout 300 python bench.py --workload highcard --steps 5 --no-e2e --no-cpu > gpurun_out/bench_highcard_nostream.json 2>/dev/null
date +%s
ls -la gpurun_out | head -60
