#!/bin/bash
# last visit of the round: the in-tree libraries and kernel cache as they will be judged — smoke, the streaming-runs and table-variant tests, one bench line
mkdir -p gpurun_out
exec > gpurun_out/final.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python -m pytest tests/test_gpu_stream.py tests/test_gpu_tpch.py -q -m gpu 2>&1 | tail -2
timeout 200 python bench.py --workload highcard --steps 5 --no-e2e --no-cpu 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); print('highcard ms/step %.2f kernel %.2f' % (d['ms_per_step'], d['roofline']['kernel_ms_per_step']))"
