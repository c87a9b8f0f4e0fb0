"""Operator protocol and partial -> final semantics of the CPU oracle, after
core/trino-main/src/test/java/io/trino/operator/TestHashAggregationOperator.java:477-560 (multi-page output, multiple
partial flushes), TestGroupByHash.java:70-250 (group ids in first-seen order, NULL group, rehash) and the hand-driving
loop of OperatorAssertion.java:62-138."""
import ctypes

import numpy as np
import pytest

import oracle_op
from oracle_op import OracleOperator
from presto_b200 import _native as N
from presto_b200 import tpch
from presto_b200.operators import Aggregation, TrinoException, plan_text, toPages
from presto_b200.spi import BIGINT, DOUBLE, Page, block_of
from util import assert_rows_equal_ignore_order, rows_of, run_oracle, sequential_bound


def q1_final_plan():
    inp = [tpch.VARCHAR1, tpch.VARCHAR1] + [BIGINT, DOUBLE] * 4 + [DOUBLE, BIGINT] * 3 + [BIGINT]
    aggs = [Aggregation("sum", DOUBLE, 2, DOUBLE), Aggregation("sum", DOUBLE, 4, DOUBLE), Aggregation("sum", DOUBLE, 6, DOUBLE),
            Aggregation("sum", DOUBLE, 8, DOUBLE), Aggregation("avg", DOUBLE, 10, DOUBLE), Aggregation("avg", DOUBLE, 12, DOUBLE),
            Aggregation("avg", DOUBLE, 14, DOUBLE), Aggregation("count", BIGINT, 16, BIGINT)]
    return plan_text(inp, step="final", groupByChannels=[0, 1], aggregations=aggs)


def test_partial_then_final_equals_single():
    total = 60_000
    shards = [[tpch.q1_page(0, 25_000, total)], [tpch.q1_page(25_000, 35_000, total)]]
    single = run_oracle(tpch.q1_plan(), shards[0] + shards[1])
    partial_pages = []
    for s in shards:
        op = OracleOperator(tpch.q1_plan("partial"))
        partial_pages += toPages(op, s)
        op.close()
    final = run_oracle(q1_final_plan(), partial_pages)
    assert_rows_equal_ignore_order(final, single, nkeys=2, rtol=sequential_bound(total))  # split sequential sums
    assert sorted(r[-1] for r in final) == sorted(r[-1] for r in single)


def q1_stage_plan(step):
    return q1_final_plan().replace("(step final)", "(step %s)" % step)


def intermediate_chain(make_operator):
    """PARTIAL per shard -> INTERMEDIATE over pairs of partial outputs -> FINAL (AggregationNode.Step: INTERMEDIATE takes and
    emits partial states, main/sql/planner/plan/AggregationNode.java:348-380; HashAggregationOperator treats it as
    step.isOutputPartial() / !step.isInputRaw())"""
    total = 90_000
    cuts = [0, 20_000, 45_000, 70_000, total]
    shards = [[tpch.q1_page(a, b - a, total)] for a, b in zip(cuts, cuts[1:])]
    partial = []
    for s in shards:
        op = make_operator(tpch.q1_plan("partial"))
        partial.append(toPages(op, s))
        op.close()
    inter = []
    for pair in (partial[0] + partial[1], partial[2] + partial[3]):
        op = make_operator(q1_stage_plan("intermediate"))
        inter += toPages(op, pair)
        op.close()
    op = make_operator(q1_stage_plan("final"))
    final = toPages(op, inter)
    op.close()
    return total, [pg for s in shards for pg in s], inter, final


def test_partial_intermediate_final_equals_single():
    from util import rows_of
    total, pages, inter, final = intermediate_chain(OracleOperator)
    single = run_oracle(tpch.q1_plan(), pages)
    assert "(step intermediate)" in q1_stage_plan("intermediate")
    assert all(len(pg.blocks) == 17 for pg in inter)           # keys + the 15 state channels: still partial states
    assert_rows_equal_ignore_order(rows_of(final), single, nkeys=2, rtol=sequential_bound(total))
    assert sorted(r[-1] for r in rows_of(final)) == sorted(r[-1] for r in single)


def test_partial_flushes_when_full_and_multi_page_output():
    # HashAggregationOperator.needsInput (:367-378) turns false once the partial table is full; getOutput then
    # flushes it in pages of <= maxOutputRows rows and the operator accepts input again.
    keys = np.arange(10_000) % 3_000
    vals = np.ones(10_000)
    pages = [Page(block_of(BIGINT, keys[i:i + 1000].tolist()), block_of(DOUBLE, vals[i:i + 1000].tolist())) for i in range(0, 10_000, 1000)]
    plan = plan_text([BIGINT, DOUBLE], step="partial", groupByChannels=[0], aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE)],
                     expectedGroups=100, maxPartialGroups=1500, maxOutputRows=400)
    op = OracleOperator(plan)
    out, refused = [], 0
    it = iter(pages)
    pending = next(it, None)
    while pending is not None:
        if op.needsInput():
            op.addInput(pending)
            pending = next(it, None)
        else:
            refused += 1
        p = op.getOutput()
        if p is not None:
            assert 0 < p.position_count <= 400
            out.append(p)
    op.finish()
    assert not op.needsInput()
    while not op.isFinished():
        p = op.getOutput()
        if p is not None:
            out.append(p)
    with pytest.raises(TrinoException):
        op.addInput(pages[0])
    op.close()
    assert refused > 0 and len(out) > 4
    rows = rows_of(out)
    assert len(rows) > 3_000          # the same key appears in several flushes
    final_plan = plan_text([BIGINT, BIGINT, DOUBLE], step="final", groupByChannels=[0], aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE)])
    final = run_oracle(final_plan, out)
    exp = {}
    for k in keys.tolist():
        exp[k] = exp.get(k, 0.0) + 1.0
    assert_rows_equal_ignore_order(final, [(k, v) for k, v in exp.items()], nkeys=1)


def _group_ids(types, blocks, expected=100):
    L = oracle_op.lib()
    plan = ("(plan (input %s))" % " ".join(str(t) for t in types)).encode()
    n = blocks[0].position_count
    cols = (N.Column * len(blocks))()
    keep = []
    for i, b in enumerate(blocks):
        b._fill(cols[i], keep)
    ids = np.zeros(n, dtype=np.int64)
    cnt = ctypes.c_int64()
    st = L.oracle_group_ids(plan, len(plan), cols, len(blocks), n, expected, ids.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)), ctypes.byref(cnt))
    assert st == 0
    return ids.tolist(), cnt.value


def test_group_ids_first_seen_order():
    # TestGroupByHash.testAddPage / testGetGroupIds: a sequence page gets ids 0..n-1; feeding it twice repeats them
    vals = list(range(100, 200))
    ids, n = _group_ids([BIGINT], [block_of(BIGINT, vals + vals)])
    assert n == 100 and ids == list(range(100)) * 2
    # testNullGroup: NULL is a group of its own, ids still in first-seen order
    ids, n = _group_ids([BIGINT], [block_of(BIGINT, [None, 5, None, 7, 5])])
    assert n == 3 and ids == [0, 1, 0, 2, 1]
    # testRehash: far more groups than expected
    vals = list(range(50_000))
    ids, n = _group_ids([BIGINT], [block_of(BIGINT, vals)], expected=4)
    assert n == 50_000 and ids == vals
    # multi channel
    ids, n = _group_ids([BIGINT, DOUBLE], [block_of(BIGINT, [1, 1, 2, 1]), block_of(DOUBLE, [0.5, 0.25, 0.5, 0.5])])
    assert n == 3 and ids == [0, 1, 2, 0]


def test_sum_bigint_overflow_raises():
    # LongSumAggregation.java:38 -> BigintOperators.add -> NUMERIC_VALUE_OUT_OF_RANGE
    plan = plan_text([BIGINT], step="single", aggregations=[Aggregation("sum", BIGINT, 0, BIGINT)])
    with pytest.raises(TrinoException) as e:
        run_oracle(plan, [Page(block_of(BIGINT, [2 ** 62, 2 ** 62]))])
    assert e.value.code == "NUMERIC_VALUE_OUT_OF_RANGE"
