"""Device side of the partial -> final exchange on one GPU: pgpu_partition_output_device places every row in the
partition the reference's function assigns, and PARTIAL -> partition -> FINAL equals the single-step result."""
import ctypes

import numpy as np
import pytest

from presto_b200 import _native as N
from presto_b200 import exchange as X
from presto_b200 import tpch
from presto_b200.operators import Aggregation, GpuPipelineOperator, plan_text
from presto_b200.spi import BIGINT, DOUBLE
from util import assert_rows_equal_ignore_order, rows_of, run_gpu, run_oracle, sequential_bound

pytestmark = pytest.mark.gpu


def _d2h(ptr, nbytes, dtype):
    out = np.empty(nbytes // np.dtype(dtype).itemsize, dtype=dtype)
    if nbytes:
        assert N.lib().pgpu_memcpy_d2h(0, out.ctypes.data, ptr, nbytes) == 0
    return out


@pytest.mark.parametrize("nparts", [1, 3, 8])
def test_partition_output_device(nparts):
    total = 300_000
    pages = [tpch.highcard_page(0, total, total, shuffled=True)]
    op = GpuPipelineOperator(tpch.highcard_plan(1000, "partial"), 0)
    try:
        for p in pages:
            op.addInput(p)
        op.finish()
        L = N.lib()
        cols = (N.Column * 8)()
        nc = ctypes.c_int32()
        offs = (ctypes.c_int64 * (nparts + 1))()
        cnts = (ctypes.c_int64 * nparts)()
        op._check(L.pgpu_partition_output_device(op._h, nparts, cols, 8, ctypes.byref(nc), offs, cnts))
        assert nc.value == 3 and op.isFinished()
        padded = offs[nparts]
        assert all(offs[i] % 16 == 0 and offs[i] + cnts[i] <= offs[i + 1] for i in range(nparts))   # aligned, disjoint
        assert all(cols[c].values % 16 == 0 and cols[c].nulls % 16 == 0 for c in range(3))
        live = np.concatenate([np.arange(offs[i], offs[i] + cnts[i]) for i in range(nparts)])
        keys_p = _d2h(cols[0].values, padded * 8, np.int64)
        keys = keys_p[live]
        cnt = _d2h(cols[1].values, padded * 8, np.int64)[live]
        sums = _d2h(cols[2].values, padded * 8, np.float64)[live]
    finally:
        op.close()
    exp = run_oracle(tpch.highcard_plan(1000, "partial"), pages)
    rows = sum(cnts)
    assert rows == len(exp) and len(set(keys.tolist())) == rows
    for part in range(nparts):
        for k in keys_p[offs[part]:offs[part] + cnts[part]][:2000].tolist():
            assert X.partition_of_row([k], [BIGINT], nparts) == part
    assert_rows_equal_ignore_order(list(zip(keys.tolist(), cnt.tolist(), sums.tolist())), exp, nkeys=1)


def test_partial_partition_final_equals_single():
    import torch
    total = 400_000
    pages = [tpch.highcard_page(0, total // 2, total), tpch.highcard_page(total // 2, total // 2, total)]
    single = run_gpu(tpch.highcard_plan(1000), pages)
    partial = GpuPipelineOperator(tpch.highcard_plan(1000, "partial"), 0)
    final = GpuPipelineOperator(plan_text([BIGINT, BIGINT, DOUBLE], step="final", groupByChannels=[0],
                                          aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE)], expectedGroups=200_000), 0)
    try:
        for p in pages:
            partial.addInput(p)
        partial.finish()
        ex = X.DeviceRepartition(0, 1, 0)
        n = ex.run(partial, final)   # world 1: partition + FINAL merge on the same device, no collective
        assert n == len(single)
        out = []
        while not final.isFinished():
            p = final.getOutput()
            if p is not None:
                out.append(p)
        assert_rows_equal_ignore_order(rows_of(out), single, nkeys=1)
    finally:
        partial.close()
        final.close()


def test_packed_partial_frames_feed_final():
    """gather shape without Python pages: PARTIAL output packed from the library's buffers, FINAL fed from the frames"""
    import numpy as np
    from presto_b200 import _native as N
    from test_oracle_protocol import q1_final_plan
    total = 300_000
    single = run_gpu(tpch.q1_plan(), [tpch.q1_page(0, total, total)])
    frames = []
    for part in range(2):
        op = GpuPipelineOperator(tpch.q1_plan("partial"), 0)
        try:
            op.addInput(tpch.q1_page(part * (total // 2), total // 2, total))
            op.finish()
            frames.append(np.frombuffer(X.pack_output(op), dtype=np.uint8).copy())
        finally:
            op.close()
    final = GpuPipelineOperator(q1_final_plan(), 0)
    try:
        assert sum(X.add_packed(final, f) for f in frames) == 8   # 4 groups per partial page
        final.finish()
        out = []
        while not final.isFinished():
            p = final.getOutput()
            if p is not None:
                out.append(p)
        assert_rows_equal_ignore_order(rows_of(out), single, nkeys=2, rtol=sequential_bound(total))
    finally:
        final.close()


def _peer_worker(rank, world, port, total, q):
    import os
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    partial = final = ex = None
    try:
        per = total // world
        partial = GpuPipelineOperator(tpch.highcard_plan(1000, "partial"), 0)
        final = GpuPipelineOperator(plan_text([BIGINT, BIGINT, DOUBLE], step="final", groupByChannels=[0],
                                              aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE)], expectedGroups=200_000), 0)
        ex = X.DeviceRepartition(0, world, rank, transport="peer")
        out = []
        for step in range(2):                                  # the second exchange reuses the cached mappings
            partial.reset()
            final.reset()
            partial.addInput(tpch.highcard_page(rank * per, per, total, shuffled=True))
            partial.finish()
            ex.run(partial, final)
            out = []
            while not final.isFinished():
                p = final.getOutput()
                if p is not None:
                    out.append(p)
        rows = rows_of(out)
        assert all(X.partition_of_row([r[0]], [BIGINT], world) == rank for r in rows[:5000])
        q.put((rank, rows, ex.sent_bytes))
    except BaseException:
        import traceback
        q.put((rank, traceback.format_exc(), -1))   # the parent fails at once instead of waiting for its queue timeout
        raise
    finally:
        if ex is not None:
            ex.close()
        for op in (partial, final):
            if op is not None:
                op.close()
        dist.destroy_process_group()


def test_peer_memory_exchange_two_processes():
    """two processes (ranks) on this GPU: each FINAL kernel reads the other process's partition buffers through CUDA IPC"""
    import os
    import torch.multiprocessing as mp
    world, total, port = 2, 200_000, 33000 + os.getpid() % 2000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_peer_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = []
    for _ in range(world):
        results.append(q.get(timeout=300))
        assert results[-1][2] >= 0, results[-1][1]
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    per = total // world
    single = run_gpu(tpch.highcard_plan(1000), [tpch.highcard_page(r * per, per, total, shuffled=True) for r in range(world)])
    got = [row for _, rows, _ in results for row in rows]
    assert all(sent > 0 for _, _, sent in results)
    assert_rows_equal_ignore_order(got, single, nkeys=1, rtol=1e-12)


def test_device_cells_feed_final():
    """PARTIAL output stays in HBM (VARCHAR keys as 16-byte cells) and is fed to FINAL as a device page"""
    from test_oracle_protocol import q1_final_plan
    total = 300_000
    single = run_gpu(tpch.q1_plan(), [tpch.q1_page(0, total, total)])
    L = N.lib()
    final = GpuPipelineOperator(q1_final_plan(), 0)
    parts = []
    try:
        for part in range(2):
            op = GpuPipelineOperator(tpch.q1_plan("partial"), 0)
            parts.append(op)
            op.addInput(tpch.q1_page(part * (total // 2), total // 2, total))
            op.finish()
            n = L.pgpu_output_channel_count(op._h)
            cols = (N.Column * n)()
            nc, rows = ctypes.c_int32(), ctypes.c_int64()
            op._check(L.pgpu_get_output_device(op._h, cols, n, ctypes.byref(nc), ctypes.byref(rows)))
            assert rows.value == 4 and cols[0].encoding == N.ENC_VARCELL and cols[1].encoding == N.ENC_VARCELL
            final.addDeviceInput(cols, nc.value, rows.value)
        final.finish()
        out = []
        while not final.isFinished():
            p = final.getOutput()
            if p is not None:
                out.append(p)
        assert_rows_equal_ignore_order(rows_of(out), single, nkeys=2, rtol=sequential_bound(total))
    finally:
        final.close()
        for op in parts:
            op.close()


def _peer_gather_worker(rank, world, port, total, q):
    import os
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import torch.distributed as dist
    from test_oracle_protocol import q1_final_plan
    dist.init_process_group("gloo", rank=rank, world_size=world)
    partial = final = ex = None
    try:
        per = total // world
        partial = GpuPipelineOperator(tpch.q1_plan("partial"), 0)
        final = GpuPipelineOperator(q1_final_plan(), 0)
        ex = X.PeerGather(0, world, rank, "pgpu_test_gather_%d" % port)
        page = tpch.q1_page(rank * per, per, total)
        rows = []
        for step in range(3):                                  # later exchanges reuse the mappings and the mailbox
            partial.reset()
            final.reset()
            partial.addInput(page)
            partial.finish()
            assert ex.run(partial, final) == 4 * world
            out = []
            while not final.isFinished():
                p = final.getOutput()
                if p is not None:
                    out.append(p)
            rows = rows_of(out)
        q.put((rank, rows, 0))
    except BaseException:
        import traceback
        q.put((rank, traceback.format_exc(), -1))
        raise
    finally:
        if ex is not None:
            ex.close()
        for op in (partial, final):
            if op is not None:
                op.close()
        dist.destroy_process_group()


def test_peer_gather_two_processes():
    """small group counts: every rank's FINAL reads both ranks' partial pages through CUDA IPC; no data-path collective"""
    import os
    import torch.multiprocessing as mp
    world, total, port = 2, 200_000, 35000 + os.getpid() % 2000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_peer_gather_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = []
    for _ in range(world):
        results.append(q.get(timeout=300))
        assert results[-1][2] >= 0, results[-1][1]
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    single = run_gpu(tpch.q1_plan(), [tpch.q1_page(0, total, total)])
    for _, rows, _ in results:
        assert_rows_equal_ignore_order(rows, single, nkeys=2, rtol=sequential_bound(total))
