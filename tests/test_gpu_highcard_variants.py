"""High-cardinality kernel variants against the oracle: run compaction through warp-private shared memory
(P::RUNS, HashCtx::consume_runs) and locality-aware table placement (PGPU_LOCB) are selected per pipeline with the
PGPU_TUNE_RUNS / PGPU_TUNE_LOCB knobs, so every case runs on the default variant AND on the alternatives: clustered and
unordered keys, a filter (dead runs), nullable keys (1- and 2-word tagless slots), table growth with deferred rows.
Group placement is never observable: results must be the same bit for bit (values are integer-valued DOUBLEs, whose
sums are exact in any order)."""
import contextlib
import os

import numpy as np
import pytest

from presto_b200 import tpch
from presto_b200.operators import Aggregation, GpuPipelineOperator, plan_text, toPages
from presto_b200.relational import compare, constant, field
from presto_b200.spi import BIGINT, DOUBLE, INTEGER, IntArrayBlock, LongArrayBlock, Page
from util import assert_rows_equal_ignore_order, rows_of, run_oracle

VARIANTS = [(0, 0), (1, 0), (0, 4), (1, 4), (1, 6)]       # (PGPU_TUNE_RUNS, PGPU_TUNE_LOCB), one key per slot (PGPU_TUNE_BUCKETS=0)
BUCKETS = "B"                                              # in place of the LOCB value: the bucketed table (PGPU_TUNE_BUCKETS=1)
NO_DIRECTORY = (1 << 20) + 1                               # expectedGroups above 2^20: no shared-memory directory, HBM table only


@contextlib.contextmanager
def variant(runs, locb):
    old = {k: os.environ.get(k) for k in ("PGPU_TUNE_RUNS", "PGPU_TUNE_LOCB", "PGPU_TUNE_BUCKETS")}
    os.environ["PGPU_TUNE_RUNS"], os.environ["PGPU_TUNE_LOCB"] = str(runs), "0" if locb == BUCKETS else str(locb)
    os.environ["PGPU_TUNE_BUCKETS"] = "1" if locb == BUCKETS else "0"
    try:
        yield
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def run_variant(plan, pages, runs, locb, device=0):
    """-> (rows, stats, generated source); device < 0 only compiles (CPU prewarm of the kernel cache)"""
    with variant(runs, locb):
        op = GpuPipelineOperator(plan, device)
        try:
            if device < 0:
                for pg in pages:
                    op.compileSignature([0] * len(pg.blocks), [b.mayHaveNull() for b in pg.blocks])
                return None, None, op.generatedSource()
            rows = rows_of(toPages(op, pages))
            return rows, op.stats(), op.generatedSource()
        finally:
            op.close()


def check_source(src, runs, locb, bucketable=True):
    assert ("RUNS = true" in src) == bool(runs)
    if locb == BUCKETS:
        assert ("BUCKETS = true" in src) == bucketable
        return
    assert "BUCKETS = false" in src
    assert ("#define PGPU_LOCB %d\n" % locb in src) == (locb > 0)


# ------------------------------------------------------------------------------------------------- cases
def highcard_case(shuffled):
    total = 400_000
    pages = [tpch.highcard_page(0, 150_001, total, shuffled), tpch.highcard_page(150_001, total - 150_001, total, shuffled)]
    return tpch.highcard_plan(NO_DIRECTORY), pages, 1


def runs_page(n, seed, key_null_frac):
    """keys in runs of 1..6 rows (some runs NULL), a second key that splits some of them, integer-valued quantities"""
    rng = np.random.default_rng(seed)
    lens = rng.integers(1, 7, n)
    k = np.repeat(rng.integers(-50_000, 50_000, n), lens)[:n]
    knull = np.repeat(rng.random(n) < key_null_frac, lens)[:n]
    k2 = np.repeat(rng.integers(0, 1 << 40, n), rng.integers(1, 4, n))[:n]
    q = rng.integers(1, 51, n).astype(np.float64)
    qnull = rng.random(n) < 0.1
    return Page(IntArrayBlock(INTEGER, k, knull.astype(np.uint8)), LongArrayBlock(BIGINT, k2), LongArrayBlock(DOUBLE, q, qnull.astype(np.uint8)))


def filtered_case(two_keys):
    """filter quantity < 30 (NULL quantity fails it: dead runs), GROUP BY a nullable INTEGER (5 key bytes: one word) or
    (BIGINT, nullable INTEGER) (13 key bytes: two words)"""
    types = [INTEGER, BIGINT, DOUBLE]
    plan = plan_text(types, compare("lt", field(2, DOUBLE), constant(30.0, DOUBLE)), [field(0, INTEGER), field(1, BIGINT), field(2, DOUBLE)],
                     step="single", groupByChannels=[1, 0] if two_keys else [0],
                     aggregations=[Aggregation("sum", DOUBLE, 2, DOUBLE), Aggregation("count", BIGINT)], expectedGroups=NO_DIRECTORY)
    pages = [runs_page(190_003, 7, 0.05), runs_page(64, 8, 0.5), runs_page(100_000, 9, 0.0)]
    return plan, pages, 2 if two_keys else 1


CASES = {"clustered": lambda: highcard_case(False), "shuffled": lambda: highcard_case(True),
         "filtered_int_key": lambda: filtered_case(False), "filtered_two_keys": lambda: filtered_case(True)}
_expected = {}


def expected_of(name):
    if name not in _expected:
        plan, pages, nkeys = CASES[name]()
        _expected[name] = (plan, pages, nkeys, run_oracle(plan, pages))
    return _expected[name]


@pytest.mark.gpu
@pytest.mark.parametrize("runs,locb", VARIANTS + [(1, BUCKETS), (0, BUCKETS)])
@pytest.mark.parametrize("name", list(CASES))
def test_variant_equals_oracle(name, runs, locb):
    plan, pages, nkeys, exp = expected_of(name)
    act, st, src = run_variant(plan, pages, runs, locb)
    check_source(src, runs, locb, bucketable=name != "filtered_two_keys")   # two key fields: no bucket of neighbouring keys
    assert st.table_bucketed == (1 if locb == BUCKETS and name != "filtered_two_keys" else 0)
    assert len(exp) > 40_000
    assert_rows_equal_ignore_order(act, exp, nkeys=nkeys, rtol=0.0)


@pytest.mark.gpu
@pytest.mark.parametrize("name,bucketed", [("clustered", 1), ("shuffled", 0), ("filtered_int_key", 0)])
def test_table_format_follows_the_key_clustering(name, bucketed):
    """PGPU_TUNE_BUCKETS=auto: the first page's keys decide (l_orderkey runs -> buckets of 8 neighbouring keys; random keys -> one
    key per slot). Without the knob the table is one key per slot (the measured default)."""
    plan, pages, nkeys, exp = expected_of(name)
    old = os.environ.pop("PGPU_TUNE_BUCKETS", None)
    op = GpuPipelineOperator(plan, 0)
    try:
        assert_rows_equal_ignore_order(rows_of(toPages(op, pages)), exp, nkeys=nkeys, rtol=0.0)
        assert op.stats().table_bucketed == 0
    finally:
        op.close()
    os.environ["PGPU_TUNE_BUCKETS"] = "auto"
    try:
        op = GpuPipelineOperator(plan, 0)
        try:
            act = rows_of(toPages(op, pages))
            assert op.stats().table_bucketed == bucketed
            op.reset()                                            # the choice survives a reset; so does the result
            again = rows_of(toPages(op, pages))
            assert op.stats().table_bucketed == bucketed
        finally:
            op.close()
    finally:
        os.environ.pop("PGPU_TUNE_BUCKETS", None)
        if old is not None:
            os.environ["PGPU_TUNE_BUCKETS"] = old
    assert_rows_equal_ignore_order(act, exp, nkeys=nkeys, rtol=0.0)
    assert_rows_equal_ignore_order(again, exp, nkeys=nkeys, rtol=0.0)


def bucket_edge_pages():
    """keys around the bucket boundaries of a signed BIGINT key: negative keys, INT64 extremes, NULL keys, all 8 keys of a
    bucket, runs that cross pages; integer-valued quantities with NULLs (a group whose quantities are all NULL still exists)"""
    rng = np.random.default_rng(21)
    base = np.array([-(1 << 63), -(1 << 63) + 7, -17, -16, -9, -8, -1, 0, 1, 7, 8, 9, 15, 16, (1 << 63) - 8, (1 << 63) - 1, 1 << 40, (1 << 40) + 3], dtype=np.int64)
    pages = []
    for n in (5000, 1, 2777, 4096):
        k = np.repeat(base[rng.integers(0, len(base), n)], rng.integers(1, 5, n))[:n]
        dense = (np.arange(n, dtype=np.int64) // 3) - 700          # clustered part: every key of many buckets, runs of 3
        k = np.where(rng.random(n) < 0.5, k, dense)
        knull = (rng.random(n) < 0.03).astype(np.uint8)
        q = rng.integers(1, 51, n).astype(np.float64)
        qnull = (rng.random(n) < 0.3).astype(np.uint8)
        pages.append(Page(LongArrayBlock(BIGINT, k, knull), LongArrayBlock(DOUBLE, q, qnull)))
    return pages


@pytest.mark.gpu
@pytest.mark.parametrize("off", ["0", "7", "3"])
@pytest.mark.parametrize("step", ["single", "partial_final"])
def test_bucketed_table_edges(step, off, monkeypatch):
    monkeypatch.setenv("PGPU_TUNE_BUCKET_OFF", off)     # buckets cut at 8n - off: keys wrap around the ends of the BIGINT range
    pages = bucket_edge_pages()
    aggs = [Aggregation("sum", DOUBLE, 1, DOUBLE), Aggregation("count", BIGINT), Aggregation("max", DOUBLE, 1, DOUBLE)]
    single = plan_text([BIGINT, DOUBLE], step="single", groupByChannels=[0], aggregations=aggs, expectedGroups=NO_DIRECTORY)
    exp = run_oracle(single, pages)
    with variant(1, BUCKETS):
        if step == "single":
            op = GpuPipelineOperator(single, 0)
            try:
                act = rows_of(toPages(op, pages))
                assert op.stats().table_bucketed == 1
            finally:
                op.close()
        else:
            partial = plan_text([BIGINT, DOUBLE], step="partial", groupByChannels=[0], aggregations=aggs, expectedGroups=NO_DIRECTORY,
                                maxPartialGroups=400)
            pop = GpuPipelineOperator(partial, 0)
            try:
                mid = toPages(pop, pages)
                assert pop.stats().table_bucketed == 1 and len(mid) >= 2
            finally:
                pop.close()
            final = plan_text([BIGINT, BIGINT, DOUBLE, BIGINT, DOUBLE], step="final", groupByChannels=[0],
                              aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE), Aggregation("count", BIGINT, 3, BIGINT), Aggregation("max", DOUBLE, 4, DOUBLE)],
                              expectedGroups=NO_DIRECTORY)
            fop = GpuPipelineOperator(final, 0)
            try:
                act = rows_of(toPages(fop, mid))
                assert fop.stats().table_bucketed == 1
            finally:
                fop.close()
    assert_rows_equal_ignore_order(act, exp, nkeys=1, rtol=0.0)


def growth_pages():
    rng = np.random.default_rng(11)
    distinct = 1_700_000
    keys = np.concatenate([rng.permutation(distinct), np.repeat(rng.permutation(distinct)[:300_000], 2)]).astype(np.int64) * 3 + 1
    q = rng.integers(1, 51, len(keys)).astype(np.float64)
    cut = [0, 900_001, 2_000_000, len(keys)]
    return keys, q, [Page(LongArrayBlock(BIGINT, keys[a:b]), LongArrayBlock(DOUBLE, q[a:b])) for a, b in zip(cut, cut[1:])]


@pytest.mark.gpu
@pytest.mark.parametrize("runs,locb", [(0, 0), (1, 0), (1, 4), (0, 4), (1, BUCKETS)])
def test_variant_table_growth(runs, locb):
    """1.7 M distinct keys into a table sized for 1.05 M: rows that need a slot past the load-factor limit are deferred,
    the table is grown and the rows are replayed (whole runs in the run-compaction variant)"""
    keys, q, pages = growth_pages()
    with variant(runs, locb):
        op = GpuPipelineOperator(tpch.highcard_plan(NO_DIRECTORY), 0)
        try:
            out = toPages(op, pages)
            st = op.stats()
            check_source(op.generatedSource(), runs, locb)
        finally:
            op.close()
    k = np.concatenate([p.blocks[0].values for p in out])
    s = np.concatenate([p.blocks[1].values.view(np.float64) for p in out])
    order = np.argsort(k, kind="stable")
    uk, inv = np.unique(keys, return_inverse=True)
    assert st.table_rehashes >= 1
    assert np.array_equal(k[order], uk)
    assert np.array_equal(s[order], np.bincount(inv, weights=q))


def prewarm():
    """compile every variant of every case into the kernel cache (no GPU needed)"""
    for name in CASES:
        plan, pages, _ = CASES[name]()
        for runs, locb in VARIANTS:
            run_variant(plan, pages, runs, locb, device=-1)
    _, _, pages = growth_pages()
    for runs, locb in VARIANTS:
        run_variant(tpch.highcard_plan(NO_DIRECTORY), pages, runs, locb, device=-1)


@pytest.mark.parametrize("runs,locb", VARIANTS + [(1, BUCKETS), (0, BUCKETS)])
def test_variants_compile_without_gpu(runs, locb):
    for name in ("clustered", "filtered_int_key", "filtered_two_keys"):
        plan, pages, _ = CASES[name]()
        _, _, src = run_variant(plan, pages[:1], runs, locb, device=-1)
        check_source(src, runs, locb, bucketable=name != "filtered_two_keys")


@pytest.mark.gpu
def test_table_format_decided_inside_a_chunked_device_page():
    """a device page longer than one launch chunk (64 M rows): the clustering probe picks the bucketed table while the first chunk
    is being launched; the remaining chunks must run the bucketed kernels too (size-independent checks: group count, sum of sums)"""
    rows = 28 * 2_500_000      # 70 M rows, 17.5 M orders
    old = os.environ.get("PGPU_TUNE_BUCKETS")
    os.environ["PGPU_TUNE_BUCKETS"] = "auto"
    dp = tpch.DevicePage(0, tpch.HC_DEVICE_SPEC, rows)
    op = GpuPipelineOperator(tpch.highcard_plan(rows // 4), 0)
    try:
        op.addDeviceInput(dp.cols, 2, rows)
        out = toPages(op, [])
        st = op.stats()
    finally:
        op.close()
        dp.free()
        os.environ.pop("PGPU_TUNE_BUCKETS", None)
        if old is not None:
            os.environ["PGPU_TUNE_BUCKETS"] = old
    assert st.table_bucketed == 1 and st.table_rehashes == 0
    k = np.concatenate([p.blocks[0].values for p in out])
    s = np.concatenate([p.blocks[1].values.view(np.float64) for p in out])
    assert len(k) == rows // 4 and len(np.unique(k)) == rows // 4
    assert st.table_slots_claimed * 8 == rows // 4                      # 8 neighbouring order keys per bucket
    qty = tpch.host_column("quantity", 0, rows, rows)
    assert float(s.sum()) == float(qty.sum())                           # integer-valued doubles: exact in any order


@pytest.mark.gpu
@pytest.mark.parametrize("runs", [0, 1])
@pytest.mark.parametrize("name", ["clustered", "shuffled", "filtered_int_key"])
def test_partial_step_variants_feed_final(name, runs):
    """the PARTIAL step with and without run compaction (its default since round 2), several partial flushes, merged by a FINAL
    step: equal to the oracle's SINGLE-step answer bit for bit"""
    plan, pages, nkeys, exp = expected_of(name)
    partial = plan.replace("(step single)", "(step partial)").rstrip().rstrip(")") + " (options (max-partial-groups 30000)))"
    with variant(runs, 0):
        pop = GpuPipelineOperator(partial, 0)
        try:
            mid = toPages(pop, pages)
            assert ("RUNS = true" in pop.generatedSource()) == bool(runs)
            assert len(mid) >= 2
        finally:
            pop.close()
    if name == "filtered_int_key":      # sum(double) -> (count, sum); count(*) -> count
        fplan = plan_text([INTEGER, BIGINT, DOUBLE, BIGINT], step="final", groupByChannels=[0],
                          aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE), Aggregation("count", BIGINT, 3, BIGINT)], expectedGroups=NO_DIRECTORY)
    else:
        fplan = plan_text([BIGINT, BIGINT, DOUBLE], step="final", groupByChannels=[0], aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE)],
                          expectedGroups=NO_DIRECTORY)
    fop = GpuPipelineOperator(fplan, 0)
    try:
        act = rows_of(toPages(fop, mid))
    finally:
        fop.close()
    assert_rows_equal_ignore_order(act, exp, nkeys=nkeys, rtol=0.0)
