"""Partial -> final exchange host logic on CPU: the partition function against a literal restatement of the reference's
formula, and the two exchange shapes run by 2 ranks over gloo with the oracle as the operator
(the N > 1 path; on GPUs the same code runs over NCCL with the CUDA operators)."""
import os
import sys

import numpy as np
import pytest

from presto_b200 import exchange as X
from presto_b200.spi import BIGINT, DATE, DOUBLE, INTEGER

M64 = (1 << 64) - 1


def _rotl(v, s):
    return ((v << s) | (v >> (64 - s))) & M64


def _hash_long(v):  # spi/type/AbstractLongType.java:126-130
    return (_rotl((v & M64) * 0xC2B2AE3D27D4EB4F & M64, 31) * 0x9E3779B185EBCA87) & M64


def _partition(values, types, n):  # HashGenerator.getPartition + InterpretedHashGenerator + CombineHashFunction
    h = 0
    for v, t in zip(values, types):
        if v is None:
            hc = 0
        elif t.kind == 4:
            bits = int(np.array([0.0 if v == 0 else v], dtype=np.float64).view(np.uint64)[0])
            if v != v:
                bits = 0x7ff8000000000000
            hc = _hash_long(bits)
        else:
            hc = _hash_long(v)
        h = (31 * h + hc) & M64
    return (h & 0x7FFFFFFFFFFFFFFF) % n


def test_partition_function_matches_reference_formula():
    rng = np.random.default_rng(7)
    keys = [0, 1, -1, 2 ** 62, -(2 ** 63), 6_000_000, 123456789012] + rng.integers(-2 ** 63, 2 ** 63 - 1, 200).tolist()
    for n in (1, 2, 3, 8, 64):
        for k in keys:
            assert X.partition_of_row([k], [BIGINT], n) == _partition([k], [BIGINT], n)
    for a, b, c in [(1, 2.5, 9000), (None, -0.0, 1), (7, float("nan"), None), (-5, 0.0, -3)]:
        for n in (2, 8):
            assert X.partition_of_row([a, b, c], [BIGINT, DOUBLE, DATE], n) == _partition([a, b, c], [BIGINT, DOUBLE, DATE], n)
    assert X.partition_of_row([-0.0], [DOUBLE], 8) == X.partition_of_row([0.0], [DOUBLE], 8)
    assert X.partition_of_row([5], [INTEGER], 8) == X.partition_of_row([5], [BIGINT], 8)  # AbstractIntType hashes the widened value


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import torch.distributed as dist
    from oracle_op import OracleOperator
    from presto_b200 import tpch
    from presto_b200.operators import Aggregation, plan_text, toPages
    from test_oracle_protocol import q1_final_plan
    from util import rows_of
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        total = 40_000
        per = total // world
        # gather shape: Q1 partial per rank -> all ranks run FINAL over all partial pages
        op = OracleOperator(tpch.q1_plan("partial"))
        partial = toPages(op, [tpch.q1_page(rank * per, per, total)])
        op.close()
        gathered = X.gather_partial_pages(partial)
        fop = OracleOperator(q1_final_plan())
        q1 = rows_of(X.final_over_pages(fop, gathered))
        fop.close()
        # repartition shape: GROUP BY orderkey partial per rank -> hash partition -> all-to-all -> FINAL per rank
        op = OracleOperator(tpch.highcard_plan(1000, "partial"))
        partial = toPages(op, [tpch.highcard_page(rank * per, per, total, shuffled=True)])
        op.close()
        mine = X.exchange_pages_host(X.partition_pages_host(partial, 1, world))
        final_plan = plan_text([BIGINT, BIGINT, DOUBLE], step="final", groupByChannels=[0], aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE)])
        fop = OracleOperator(final_plan)
        hc = rows_of(X.final_over_pages(fop, mine))
        fop.close()
        assert all(X.partition_of_row([r[0]], [BIGINT], world) == rank for r in hc)
        q.put((rank, q1, hc))
    finally:
        dist.destroy_process_group()


def test_two_rank_exchange_over_gloo():
    import torch.multiprocessing as mp
    from presto_b200 import tpch
    from util import assert_rows_equal_ignore_order, run_oracle, sequential_bound
    world, port = 2, 29000 + os.getpid() % 2000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    total = 40_000
    single_q1 = run_oracle(tpch.q1_plan(), [tpch.q1_page(0, total, total)])
    single_hc = run_oracle(tpch.highcard_plan(1000), [tpch.highcard_page(0, total // 2, total, True), tpch.highcard_page(total // 2, total // 2, total, True)])
    hc_all = []
    for rank, q1, hc in results:
        assert_rows_equal_ignore_order(q1, single_q1, nkeys=2, rtol=sequential_bound(total))  # every rank holds the final Q1 answer
        hc_all += hc
    assert_rows_equal_ignore_order(hc_all, single_hc, nkeys=1, rtol=1e-12)  # the union of the partitions is the answer


def _frame_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import numpy as np
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ok = True
        for size in (0, 100, 40_000 + rank * 70_000):   # empty, one frame, ragged frames larger than the fixed frame
            mine = (np.arange(size, dtype=np.int64) * (rank + 3) % 251).astype(np.uint8).tobytes()
            frames = X.all_gather_frames(mine)
            for r, f in enumerate(frames):
                n = size if size < 40_000 else 40_000 + r * 70_000
                exp = (np.arange(n, dtype=np.int64) * (r + 3) % 251).astype(np.uint8)
                ok = ok and f.size == n and bool((f == exp).all())
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


def test_all_gather_frames_over_gloo():
    import torch.multiprocessing as mp
    world, port = 2, 31000 + os.getpid() % 2000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_frame_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert all(ok for _, ok in results)
