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


FINAL_HC = lambda groups: plan_text([BIGINT, BIGINT, DOUBLE], step="final", groupByChannels=[0],  # noqa: E731
                                    aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE)], expectedGroups=groups)
REPARTITION = 200_000   # expected-groups above what an inbox slot holds: pgpu_exchange_partial takes the repartition shape


def drain(op):
    out = []
    while not op.isFinished():
        p = op.getOutput()
        if p is not None and p.position_count:
            out.append(p)
    return rows_of(out)


def _unique(tag):
    import os
    import time
    return "pgpu_test_%s_%d_%d" % (tag, os.getpid(), time.time_ns() % 1_000_000_000)


def test_exchange_one_rank_repartition_equals_single(monkeypatch):
    monkeypatch.setenv("PGPU_TUNE_RANGE_SPLIT", "0")       # the plain repartition shape (the range-aware split is tested below)
    total = 400_000
    pages = [tpch.highcard_page(0, total // 2, total), tpch.highcard_page(total // 2, total // 2, total)]
    single = run_gpu(tpch.highcard_plan(1000), pages)
    partial = GpuPipelineOperator(tpch.highcard_plan(REPARTITION, "partial"), 0)
    final = GpuPipelineOperator(FINAL_HC(REPARTITION), 0)
    comm = X.Comm(1, 0, 0, _unique("one"))
    try:
        for step in range(2):
            partial.reset()
            final.reset()
            for p in pages:
                partial.addInput(p)
            partial.finish()
            assert comm.exchange(partial, final) == len(single)   # world 1: partition + FINAL merge on the same device
            assert partial.isFinished()
            assert_rows_equal_ignore_order(drain(final), single, nkeys=1)
    finally:
        comm.close()
        partial.close()
        final.close()


def test_exchange_one_rank_gather_q1_and_q6():
    """gather shape with a single rank: the push kernel stores into the rank's own inbox, the gather kernel merges the slot"""
    from test_oracle_protocol import q1_final_plan
    total = 300_000
    comm = X.Comm(1, 0, 0, _unique("g1"))
    try:
        for pplan, fplan, page_fn, nkeys in ((tpch.q1_plan("partial"), q1_final_plan(), tpch.q1_page, 2),
                                             (tpch.q6_plan("partial"), plan_text([BIGINT, DOUBLE], step="final", aggregations=[Aggregation("sum", DOUBLE, 0, DOUBLE)]), tpch.q6_page, None)):
            pages = [page_fn(0, total, total)]
            single = run_gpu(pplan.replace("(step partial)", "(step single)"), pages)
            partial, final = GpuPipelineOperator(pplan, 0), GpuPipelineOperator(fplan, 0)
            try:
                for step in range(3):                              # both inbox parities, then the first one again
                    partial.reset()
                    final.reset()
                    partial.addInput(pages[0])
                    partial.finish()
                    assert comm.exchange(partial, final) == -1     # row count known on the device only
                    assert_rows_equal_ignore_order(drain(final), single, nkeys=nkeys, rtol=sequential_bound(total))
            finally:
                partial.close()
                final.close()
    finally:
        comm.close()


def test_exchange_gather_overflow_is_an_error():
    """a plan that promises few groups but produces more than an inbox slot holds fails loudly at getOutput"""
    from presto_b200.operators import TrinoException
    total = 100_000
    partial = GpuPipelineOperator(tpch.highcard_plan(1000, "partial"), 0)    # expected-groups 1000: gather shape; 25 000 groups arrive
    final = GpuPipelineOperator(FINAL_HC(1000), 0)
    comm = X.Comm(1, 0, 0, _unique("ovf"))
    try:
        partial.addInput(tpch.highcard_page(0, total, total))
        partial.finish()
        comm.exchange(partial, final)
        with pytest.raises(TrinoException) as e:
            drain(final)
        assert "inbox slot" in str(e.value)
    finally:
        comm.close()
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


def _repartition_worker(rank, world, name, total, q):
    partial = final = comm = None
    try:
        per = total // world
        partial = GpuPipelineOperator(tpch.highcard_plan(REPARTITION, "partial"), 0)
        final = GpuPipelineOperator(FINAL_HC(REPARTITION), 0)
        comm = X.Comm(world, rank, 0, name)
        rows = []
        for step in range(3):                                  # later exchanges reuse the mapped arenas
            partial.reset()
            final.reset()
            partial.addInput(tpch.highcard_page(rank * per, per, total, shuffled=True))
            partial.finish()
            comm.exchange(partial, final)
            rows = drain(final)
        assert all(X.partition_of_row([r[0]], [BIGINT], world) == rank for r in rows[:5000])
        q.put((rank, rows, comm.sent_bytes))
    except BaseException:
        import traceback
        q.put((rank, traceback.format_exc(), -1))   # the parent fails at once instead of waiting for its queue timeout
        raise
    finally:
        if comm is not None:
            comm.close()
        for op in (partial, final):
            if op is not None:
                op.close()


def _spawn(worker, world, name, total):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=worker, args=(r, world, name, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = []
    for _ in range(world):
        results.append(q.get(timeout=300))
        assert results[-1][2] >= 0, results[-1][1]
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    return results


def test_exchange_repartition_two_processes():
    """two processes (ranks) on this GPU: each FINAL kernel reads the other process's partition arena through CUDA IPC"""
    world, total = 2, 200_000
    results = _spawn(_repartition_worker, world, _unique("rp"), total)
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


def _gather_worker(rank, world, name, total, q):
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from test_oracle_protocol import q1_final_plan
    partial = final = comm = None
    try:
        per = total // world
        partial = GpuPipelineOperator(tpch.q1_plan("partial"), 0)
        final = GpuPipelineOperator(q1_final_plan(), 0)
        comm = X.Comm(world, rank, 0, name)
        page = tpch.q1_page(rank * per, per, total)
        rows = []
        for step in range(4):                                  # both inbox parities twice
            partial.reset()
            final.reset()
            partial.addInput(page)
            partial.finish()
            comm.exchange(partial, final)
            rows = drain(final)
        q.put((rank, rows, 0))
    except BaseException:
        import traceback
        q.put((rank, traceback.format_exc(), -1))
        raise
    finally:
        if comm is not None:
            comm.close()
        for op in (partial, final):
            if op is not None:
                op.close()


def test_exchange_gather_two_processes():
    """small group counts: each rank's push kernel stores its partial rows into both ranks' inboxes (CUDA IPC), each rank's
    gather kernel merges both slots; no host round trip, no collective"""
    world, total = 2, 200_000
    results = _spawn(_gather_worker, world, _unique("ga"), total)
    single = run_gpu(tpch.q1_plan(), [tpch.q1_page(0, total, total)])
    for _, rows, _ in results:
        assert_rows_equal_ignore_order(rows, single, nkeys=2, rtol=sequential_bound(total))


# ---- range-aware split of the repartition shape -----------------------------------------------------------------------------
def test_range_split_one_rank_finalises_everything_in_place():
    """world 1: no other rank holds any key, so every group is private: no partition pass, no FINAL hash build"""
    total = 400_000
    pages = [tpch.highcard_page(0, total // 2, total), tpch.highcard_page(total // 2, total // 2, total)]
    single = run_gpu(tpch.highcard_plan(1000), pages)
    partial = GpuPipelineOperator(tpch.highcard_plan(REPARTITION, "partial"), 0)
    final = GpuPipelineOperator(FINAL_HC(REPARTITION), 0)
    comm = X.Comm(1, 0, 0, _unique("rs1"))
    try:
        for step in range(2):
            partial.reset()
            final.reset()
            for p in pages:
                partial.addInput(p)
            partial.finish()
            assert comm.exchange(partial, final) == 0               # nothing went through the exchange
            assert comm.private_groups == len(single)
            assert final.stats().table_capacity == 0                # FINAL never built a table
            assert_rows_equal_ignore_order(drain(final), single, nkeys=1)
    finally:
        comm.close()
        partial.close()
        final.close()


def _range_split_worker(rank, world, name, total, q):
    """arg `total` < 0: shuffled keys (every rank's range covers the others': nearly everything is shared)"""
    shuffled = total < 0
    total = abs(total)
    partial = final = comm = None
    try:
        per = total // world
        partial = GpuPipelineOperator(tpch.highcard_plan(REPARTITION, "partial"), 0)
        final = GpuPipelineOperator(FINAL_HC(REPARTITION), 0)
        comm = X.Comm(world, rank, 0, name)
        rows = []
        for step in range(3):
            partial.reset()
            final.reset()
            # a cut in the middle of an order: the boundary order's rows lie on both ranks
            partial.addInput(tpch.highcard_page(rank * per, per, total, shuffled=shuffled))
            partial.finish()
            comm.exchange(partial, final)
            rows = drain(final)
        q.put((rank, rows, comm.private_groups))
    except BaseException:
        import traceback
        q.put((rank, traceback.format_exc(), -1))
        raise
    finally:
        if comm is not None:
            comm.close()
        for op in (partial, final):
            if op is not None:
                op.close()


@pytest.mark.parametrize("shuffled", [False, True])
def test_range_split_two_processes(shuffled):
    """clustered keys: each rank finalises its own orders in place, only the orders around the cut are exchanged;
    shuffled keys: the ranges overlap, (nearly) everything takes the exchange. Either way every group comes out exactly once."""
    world, total = 2, 200_002                                      # 100 001 rows per rank: the cut falls inside a 28-row block
    results = _spawn(_range_split_worker, world, _unique("rs2"), -total if shuffled else total)
    per = total // world
    single = run_gpu(tpch.highcard_plan(1000), [tpch.highcard_page(r * per, per, total, shuffled=shuffled) for r in range(world)])
    got = [row for _, rows, _ in results for row in rows]
    assert_rows_equal_ignore_order(got, single, nkeys=1, rtol=1e-12)   # exactly once each: a duplicate group would fail the pairing
    private = [p for _, _, p in results]
    if shuffled:
        assert sum(private) < len(single) // 100
    else:
        assert all(p > 0 for p in private) and sum(private) >= len(single) - 16


def _null_key_worker(rank, world, name, total, q):
    from presto_b200.spi import INTEGER, IntArrayBlock
    partial = final = comm = None
    try:
        aggs = [Aggregation("sum", DOUBLE, 1, DOUBLE), Aggregation("count", BIGINT), Aggregation("max", DOUBLE, 1, DOUBLE)]
        pplan = plan_text([INTEGER, DOUBLE], step="partial", groupByChannels=[0], aggregations=aggs, expectedGroups=REPARTITION)
        fplan = plan_text([INTEGER, BIGINT, DOUBLE, BIGINT, DOUBLE], step="final", groupByChannels=[0],
                          aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE), Aggregation("count", BIGINT, 3, BIGINT), Aggregation("max", DOUBLE, 4, DOUBLE)],
                          expectedGroups=REPARTITION)
        partial, final = GpuPipelineOperator(pplan, 0), GpuPipelineOperator(fplan, 0)
        comm = X.Comm(world, rank, 0, name)
        partial.addInput(_null_key_page(rank, total))
        partial.finish()
        comm.exchange(partial, final)
        q.put((rank, drain(final), comm.private_groups))
    except BaseException:
        import traceback
        q.put((rank, traceback.format_exc(), -1))
        raise
    finally:
        if comm is not None:
            comm.close()
        for op in (partial, final):
            if op is not None:
                op.close()


def _null_key_page(rank, n):
    """nullable INTEGER key (sign-extended 32-bit field behind a null byte): rank 0 holds negative keys up to 10, rank 1 keys from
    -5 upwards, both hold NULL keys: [-5, 10] and NULL are shared, the rest is private"""
    from presto_b200.spi import INTEGER, IntArrayBlock, Page, LongArrayBlock
    rng = np.random.default_rng(40 + rank)
    k = rng.integers(-3000, 11, n) if rank == 0 else rng.integers(-5, 4000, n)
    knull = (rng.random(n) < 0.05).astype(np.uint8)
    v = rng.integers(1, 100, n).astype(np.float64)
    vnull = (rng.random(n) < 0.2).astype(np.uint8)
    return Page(IntArrayBlock(INTEGER, k, knull), LongArrayBlock(DOUBLE, v, vnull))


def test_range_split_null_and_negative_keys():
    from presto_b200.spi import INTEGER
    world, n = 2, 60_000
    results = _spawn(_null_key_worker, world, _unique("rs3"), n)
    aggs = [Aggregation("sum", DOUBLE, 1, DOUBLE), Aggregation("count", BIGINT), Aggregation("max", DOUBLE, 1, DOUBLE)]
    single = plan_text([INTEGER, DOUBLE], step="single", groupByChannels=[0], aggregations=aggs, expectedGroups=REPARTITION)
    exp = run_oracle(single, [_null_key_page(r, n) for r in range(world)])
    got = [row for _, rows, _ in results for row in rows]
    assert_rows_equal_ignore_order(got, exp, nkeys=1, rtol=0.0)       # integer-valued doubles: exact
    assert sum(1 for r in got if r[0] is None) == 1                    # the NULL group of both ranks was merged, not emitted twice
    private = dict((rank, p) for rank, _, p in results)
    assert private[0] > 2000 and private[1] > 3000                     # keys below -5 / above 10


def test_range_split_can_be_turned_off(monkeypatch):
    monkeypatch.setenv("PGPU_TUNE_RANGE_SPLIT", "0")
    total = 200_000
    pages = [tpch.highcard_page(0, total, total)]
    single = run_gpu(tpch.highcard_plan(1000), pages)
    partial = GpuPipelineOperator(tpch.highcard_plan(REPARTITION, "partial"), 0)
    final = GpuPipelineOperator(FINAL_HC(REPARTITION), 0)
    comm = X.Comm(1, 0, 0, _unique("rs0"))
    try:
        partial.addInput(pages[0])
        partial.finish()
        assert comm.exchange(partial, final) == len(single) and comm.private_groups == 0
        assert_rows_equal_ignore_order(drain(final), single, nkeys=1)
    finally:
        comm.close()
        partial.close()
        final.close()
