#!/usr/bin/env python
"""Per-phase wall clock of one Q1 step with the native gather exchange, under torchrun (diagnostic):
  torchrun --nproc-per-node N tools/exchange_trace.py [rows]"""
import os, sys, time, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as td
import bench as B
from presto_b200 import tpch, exchange as X, _native as N
from presto_b200.operators import GpuPipelineOperator
rank, world, dev = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(dev)
td.init_process_group("nccl", device_id=torch.device("cuda", dev))
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 600_037_902
name = sys.argv[2] if len(sys.argv) > 2 else "q1"     # q1 | q6 | highcard | highcard_shuffled; PGPU_TRACE_EXCHANGE=1 adds the library's phases
wl = B.workload_def(name, rows)
dp = tpch.DevicePage(dev, wl["spec"], rows, first_row=rank * rows, total_rows=rows * world)
op = GpuPipelineOperator(wl["plan"]("partial"), dev)
fin = GpuPipelineOperator(B.final_plan_for(wl, None), dev)
comm = B.get_comm(dev, world, rank)
acc = {}
def lap(name, t0):
    t1 = time.perf_counter(); acc[name] = acc.get(name, 0.0) + t1 - t0; return t1
for it in range(13):
    if it == 3:
        acc.clear(); td.barrier(); torch.cuda.synchronize()
    t = time.perf_counter()
    op.reset(); t = lap("partial.reset", t)
    op.addDeviceInput(dp.cols, len(wl["spec"]), rows); op.finish(); t = lap("enqueue partial", t)
    if os.environ.get("TRACE_SYNC"):
        op.sync(); t = lap("partial kernels (sync)", t)
    fin.reset(); t = lap("final.reset", t)
    comm.exchange(op, fin); t = lap("exchange (enqueue)", t)
    n = B.drain_output(fin, name in ("q1", "q6")); t = lap("final drain (sync + output)", t)
td.barrier()
if rank == 0:
    print("N=%d rows/rank=%d result rows %d; ms per step:" % (world, rows, n), {k: round(v * 100, 3) for k, v in acc.items()}, "total", round(sum(acc.values()) * 100, 3))
comm.close(); td.destroy_process_group()
