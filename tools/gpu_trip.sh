#!/bin/bash
mkdir -p gpurun_out
exec > gpurun_out/trip.log 2>&1
timeout 600 python -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout 180 > gpurun_out/trip_pytest.log 2>&1; tail -3 gpurun_out/trip_pytest.log
run() {
  w=$1; shift
  env "$@" timeout 150 python bench.py --workload $w --steps 5 --no-cpu --no-e2e 2>gpurun_out/trip_$w.err | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$w $*', d['ms_per_step'], d['roofline']['frac'], d['roofline']['avg_launch_ms'], d['gpu_launches'], d['result_rows'])"
  tail -3 gpurun_out/trip_$w.err
}
run q1dict A=1
run q1 A=1
run highcard A=1
run highcard_shuffled A=1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:pgpu_k_page_staged -s 2 -c 1 -f -o gpurun_out/q1dict_d \
    python bench.py --workload q1dict --steps 1 --warmup 1 --no-e2e --no-cpu --rows 200000000 2>&1 | tail -1 | cut -c1-100
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_hc.csv \
  python bench.py --workload highcard --steps 1 --warmup 1 --no-e2e --no-cpu > /dev/null 2>&1; echo "launch list exit $?"
