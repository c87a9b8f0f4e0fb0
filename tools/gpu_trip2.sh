#!/bin/bash
# multi-GPU visit: bash tools/gpu_trip2.sh N   (run under gpurun --gpus N)
N=${1:-2}
mkdir -p gpurun_out
exec > gpurun_out/trip${N}.log 2>&1
nvidia-smi -L | head -8
run() {
  tag=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" 2>gpurun_out/scale${N}_$tag.err > gpurun_out/scale${N}_$tag.json
  echo "$tag exit $?"; tail -1 gpurun_out/scale${N}_$tag.json | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$tag', 'ms/step', round(d['ms_per_step'],3), 'rows/s %.4g' % d['value'], 'e2e', d['e2e'] and '%.4g' % d['e2e']['value'], 'result_rows', d['result_rows'], d['exchange'])"
  grep -v "^\*\*\*\|OMP_NUM\|^$" gpurun_out/scale${N}_$tag.err | tail -4
}
if [ "$2" = "trace" ]; then
  PGPU_BENCH_TRACE=1 run q1t --workload q1 --steps 20 --warmup 3 --no-cpu --no-e2e
  run q1 --workload q1 --steps 20 --warmup 3 --no-cpu --no-e2e
  run q1_nccl --workload q1 --steps 20 --warmup 3 --no-cpu --no-e2e --exchange nccl
  run q6 --workload q6 --steps 20 --warmup 3 --no-cpu --no-e2e
  exit 0
fi
if [ "$2" = "lean" ]; then
  run q1 --workload q1 --steps 10 --warmup 3 --no-cpu
  run q6 --workload q6 --steps 10 --warmup 3 --no-cpu --no-e2e
  run q1dict --workload q1dict --steps 10 --warmup 3 --no-cpu --no-e2e
  run hc_peer --workload highcard --steps 5 --warmup 3 --no-cpu --no-e2e --exchange peer
  # config 5: sf1000 (5 999 989 709 rows) split over the N GPUs, device resident
  R=$((5999989709 / N))
  if [ $R -le 1600000000 ]; then
    run q1_sf1000 --workload q1 --rows $R --steps 5 --warmup 3 --no-cpu --no-e2e
    run hc_sf1000 --workload highcard --rows $R --steps 3 --warmup 3 --no-cpu --no-e2e --exchange peer
  fi
  exit 0
fi
run q1 --workload q1 --steps 10 --warmup 3 --no-cpu
run q1_nccl --workload q1 --steps 10 --warmup 3 --no-cpu --no-e2e --exchange nccl
run q6 --workload q6 --steps 10 --warmup 3 --no-cpu --no-e2e
run q1dict --workload q1dict --steps 10 --warmup 3 --no-cpu --no-e2e
run hc_peer --workload highcard --steps 5 --warmup 3 --no-cpu --no-e2e --exchange peer
run hc_nccl --workload highcard --steps 5 --warmup 3 --no-cpu --no-e2e --exchange nccl
run hcs_peer --workload highcard_shuffled --steps 5 --warmup 3 --no-cpu --no-e2e --exchange peer
