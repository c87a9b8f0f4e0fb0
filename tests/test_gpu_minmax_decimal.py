"""min / max over long DECIMAL (AbstractMinMaxAggregationFunction.java:274-306 on Int128 values; the closed forms of
TestLongDecimalMinAggregation / TestLongDecimalMaxAggregation are in tests/kat_cases.py): a 128-bit cell updated with a 16-byte
compare-and-swap loop (csrc/device/pgpu_device.cuh max128_*). Every aggregation path — global, shared-memory directory, HBM
table with 1- and 2-word tagless keys and with tagged 3-word keys, PARTIAL -> FINAL — against the oracle and against Python's
exact integers, bit-exact. Values span the whole DECIMAL(38, s) range on both sides of zero."""
import numpy as np
import pytest

from presto_b200.operators import Aggregation, GpuPipelineOperator, plan_text, toPages
from presto_b200.spi import BIGINT, INTEGER, Int128ArrayBlock, IntArrayBlock, LongArrayBlock, Page, createDecimalType
from util import assert_rows_equal_ignore_order, rows_of, run_gpu, run_oracle

pytestmark = pytest.mark.gpu
D38 = createDecimalType(38, 4)
D20 = createDecimalType(20, 0)
NO_DIRECTORY = (1 << 20) + 1


def values(n, seed, t):
    """magnitudes of every size up to the type's bound, both signs, zeros, duplicates of the extremes"""
    rng = np.random.default_rng(seed)
    bound = 10**t.precision - 1
    digits = rng.integers(0, t.precision + 1, n)
    v = [int(rng.integers(0, 10**9)) * 10**max(int(d) - 9, 0) % (bound + 1) for d in digits]
    v = [x if rng.random() < 0.5 else -x for x in v]
    for k in range(0, n, 97):
        v[k] = bound if (k // 97) % 2 else -bound
    nulls = (rng.random(n) < 0.15).astype(np.uint8)
    return v, nulls


def keyed_page(n, ngroups, seed, t, key_type=BIGINT):
    rng = np.random.default_rng(seed + 1)
    v, nulls = values(n, seed, t)
    k = rng.integers(0, ngroups, n)
    nulls[k == 3] = 1                      # one group whose values are all NULL
    kb = LongArrayBlock(BIGINT, k) if key_type == BIGINT else IntArrayBlock(INTEGER, k.astype(np.int32))
    return Page(kb, Int128ArrayBlock(t, v, nulls)), k, v, nulls


def exact(k, v, nulls):
    out = {}
    for key, val, isnull in zip(k.tolist(), v, nulls.tolist()):
        lo, hi, cnt = out.get(key, (None, None, 0))
        if not isnull:
            lo = val if lo is None else min(lo, val)
            hi = val if hi is None else max(hi, val)
        out[key] = (lo, hi, cnt + 1)
    return sorted((key, lo, hi, cnt) for key, (lo, hi, cnt) in out.items())


def plan_for(t, expected_groups, key_type=BIGINT, step="single"):
    return plan_text([key_type, t], step=step, groupByChannels=[0],
                     aggregations=[Aggregation("min", t, 1, t), Aggregation("max", t, 1, t), Aggregation("count", BIGINT)], expectedGroups=expected_groups)


@pytest.mark.parametrize("t", [D38, D20], ids=["decimal38", "decimal20"])
def test_global_min_max(t):
    v, nulls = values(50_001, 5, t)
    page = Page(Int128ArrayBlock(t, v, nulls))
    plan = plan_text([t], step="single", aggregations=[Aggregation("min", t, 0, t), Aggregation("max", t, 0, t), Aggregation("count", BIGINT, 0, t)])
    got = run_gpu(plan, [page])
    live = [x for x, isnull in zip(v, nulls) if not isnull]
    assert got == [(min(live), max(live), len(live))]
    assert got == run_oracle(plan, [page])
    # no value at all: NULL
    empty = Page(Int128ArrayBlock(t, [1, 2, 3], np.ones(3, dtype=np.uint8)))
    assert run_gpu(plan, [empty]) == [(None, None, 0)] == run_oracle(plan, [empty])


@pytest.mark.parametrize("key_type", [BIGINT, INTEGER], ids=["two_key_words", "one_key_word"])
@pytest.mark.parametrize("expected_groups,ngroups", [(100, 40), (NO_DIRECTORY, 30_000)], ids=["directory", "hbm_table"])
def test_grouped_min_max(expected_groups, ngroups, key_type):
    page, k, v, nulls = keyed_page(200_003, ngroups, 11, D38, key_type)
    plan = plan_for(D38, expected_groups, key_type)
    got = run_gpu(plan, [page])
    assert sorted(got) == exact(k, v, nulls)
    assert_rows_equal_ignore_order(got, run_oracle(plan, [page]), nkeys=1, rtol=0.0)


def test_grouped_min_max_tagged_slots():
    """three key fields = 25 key bytes: slots carry a tag word, the 128-bit cell still has to land on a 16-byte boundary"""
    rng = np.random.default_rng(2)
    n = 60_000
    v, nulls = values(n, 21, D38)
    k1, k2, k3 = rng.integers(0, 50, n), rng.integers(-5, 5, n), rng.integers(0, 3, n)
    page = Page(LongArrayBlock(BIGINT, k1), LongArrayBlock(BIGINT, k2), LongArrayBlock(BIGINT, k3), Int128ArrayBlock(D38, v, nulls))
    for extra in ([], [Aggregation("count", BIGINT)]):   # an odd and an even number of words in front of the pair
        plan = plan_text([BIGINT, BIGINT, BIGINT, D38], step="single", groupByChannels=[0, 1, 2],
                         aggregations=extra + [Aggregation("max", D38, 3, D38), Aggregation("min", D38, 3, D38)], expectedGroups=NO_DIRECTORY)
        assert_rows_equal_ignore_order(run_gpu(plan, [page]), run_oracle(plan, [page]), nkeys=3, rtol=0.0)


def test_partial_final_min_max():
    """the intermediate state of min / max is the value itself (an Int128 block): PARTIAL pages from two pipelines merged by FINAL"""
    pages, ks, vs, ns = [], [], [], []
    for seed in (31, 32, 33):
        pg, k, v, nulls = keyed_page(70_000, 5_000, seed, D38)
        pages.append(pg)
        ks.append(k)
        vs += v
        ns.append(nulls)
    mids = []
    for part in (pages[:2], pages[2:]):
        op = GpuPipelineOperator(plan_for(D38, NO_DIRECTORY, step="partial"), 0)
        try:
            mids += toPages(op, part)
        finally:
            op.close()
    final = plan_text([BIGINT, D38, D38, BIGINT], step="final", groupByChannels=[0],
                      aggregations=[Aggregation("min", D38, 1, D38), Aggregation("max", D38, 2, D38), Aggregation("count", BIGINT, 3, BIGINT)], expectedGroups=NO_DIRECTORY)
    fop = GpuPipelineOperator(final, 0)
    try:
        got = rows_of(toPages(fop, mids))
    finally:
        fop.close()
    assert sorted(got) == exact(np.concatenate(ks), vs, np.concatenate(ns))


def test_min_max_next_to_sums_and_short_decimals():
    d12 = createDecimalType(12, 2)
    rng = np.random.default_rng(8)
    n = 90_000
    v, nulls = values(n, 41, D20)
    s = rng.integers(-10**11, 10**11, n)
    page = Page(LongArrayBlock(BIGINT, rng.integers(0, 2_000, n)), Int128ArrayBlock(D20, v, nulls), LongArrayBlock(d12, s))
    plan = plan_text([BIGINT, D20, d12], step="single", groupByChannels=[0],
                     aggregations=[Aggregation("sum", createDecimalType(38, 0), 1, D20), Aggregation("max", D20, 1, D20), Aggregation("min", d12, 2, d12),
                                   Aggregation("avg", D20, 1, D20), Aggregation("min", D20, 1, D20), Aggregation("max", d12, 2, d12)], expectedGroups=NO_DIRECTORY)
    assert_rows_equal_ignore_order(run_gpu(plan, [page]), run_oracle(plan, [page]), nkeys=1, rtol=0.0)
