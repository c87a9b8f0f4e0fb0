#!/bin/bash
mkdir -p gpurun_out
exec > gpurun_out/trip2.log 2>&1
nvidia-smi -L
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
for w in q1 highcard; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload $w --steps 5 --warmup 3 2>gpurun_out/trip2_$w.err > gpurun_out/trip2_$w.json
  echo "exit $?"; tail -1 gpurun_out/trip2_$w.json | cut -c1-1800; tail -5 gpurun_out/trip2_$w.err
done
