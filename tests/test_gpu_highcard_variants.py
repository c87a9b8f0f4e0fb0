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

VARIANTS = [(0, 0), (1, 0), (0, 4), (1, 4), (1, 6)]       # (PGPU_TUNE_RUNS, PGPU_TUNE_LOCB)
NO_DIRECTORY = (1 << 20) + 1                               # expectedGroups above 2^20: no shared-memory directory, HBM table only


@contextlib.contextmanager
def variant(runs, locb):
    old = {k: os.environ.get(k) for k in ("PGPU_TUNE_RUNS", "PGPU_TUNE_LOCB")}
    os.environ["PGPU_TUNE_RUNS"], os.environ["PGPU_TUNE_LOCB"] = str(runs), str(locb)
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


def check_source(src, runs, locb):
    assert ("RUNS = true" in src) == bool(runs)
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
@pytest.mark.parametrize("runs,locb", VARIANTS)
@pytest.mark.parametrize("name", list(CASES))
def test_variant_equals_oracle(name, runs, locb):
    plan, pages, nkeys, exp = expected_of(name)
    act, _, src = run_variant(plan, pages, runs, locb)
    check_source(src, runs, locb)
    assert len(exp) > 40_000
    assert_rows_equal_ignore_order(act, exp, nkeys=nkeys, rtol=0.0)


def growth_pages():
    rng = np.random.default_rng(11)
    distinct = 1_700_000
    keys = np.concatenate([rng.permutation(distinct), np.repeat(rng.permutation(distinct)[:300_000], 2)]).astype(np.int64) * 3 + 1
    q = rng.integers(1, 51, len(keys)).astype(np.float64)
    cut = [0, 900_001, 2_000_000, len(keys)]
    return keys, q, [Page(LongArrayBlock(BIGINT, keys[a:b]), LongArrayBlock(DOUBLE, q[a:b])) for a, b in zip(cut, cut[1:])]


@pytest.mark.gpu
@pytest.mark.parametrize("runs,locb", [(0, 0), (1, 0), (1, 4), (0, 4)])
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


@pytest.mark.parametrize("runs,locb", VARIANTS)
def test_variants_compile_without_gpu(runs, locb):
    for name in ("clustered", "filtered_int_key", "filtered_two_keys"):
        plan, pages, _ = CASES[name]()
        _, _, src = run_variant(plan, pages[:1], runs, locb, device=-1)
        check_source(src, runs, locb)
