"""Partition sweep (DESIGN.md "partition sweep"; csrc/device/pgpu_device.cuh body_part_count / body_part_scatter /
body_agg_hash_sweep): long device-resident pages are reordered by table slice before the hash-aggregation kernel runs.
Where a group's slot lies and in which order rows reach it is never observable: every case must equal the CPU oracle bit for
bit (quantities are integer-valued DOUBLEs, whose sums are exact in any order). The thresholds that keep short pages and small
tables on the ordinary path are lowered with the PGPU_TUNE_SWEEP_* knobs so that the cases stay small."""
import ctypes

import numpy as np
import pytest

from presto_b200 import _native as N
from presto_b200 import tpch
from presto_b200.operators import Aggregation, GpuPipelineOperator, plan_text, toPages
from presto_b200.relational import compare, constant, field
from presto_b200.spi import (BIGINT, DOUBLE, INTEGER, DictionaryBlock, Int128ArrayBlock, IntArrayBlock, LongArrayBlock, Page, RunLengthEncodedBlock,
                             createDecimalType)
from util import assert_rows_equal_ignore_order, rows_of, run_oracle

NO_DIRECTORY = (1 << 20) + 1     # expectedGroups above 2^20: HBM table only


class DevicePageOf:
    """a host Page uploaded as it is: the same pgpu_column array with device pointers (plain, dictionary and RLE blocks)"""

    def __init__(self, page: Page, device: int = 0):
        self.device, self.rows, self.ncols = device, page.position_count, len(page.blocks)
        self._allocs, self._keep = [], []
        self.cols, keep = GpuPipelineOperator._marshal(page.blocks)
        self._keep.append(keep)
        for i, b in enumerate(page.blocks):
            self._swap(self.cols[i], b)

    def _up(self, arr):
        arr = np.ascontiguousarray(arr)
        p = N.lib().pgpu_device_alloc(self.device, max(arr.nbytes, 16))
        assert p
        self._allocs.append(p)
        if arr.nbytes:
            assert N.lib().pgpu_memcpy_h2d(self.device, p, arr.ctypes.data, arr.nbytes) == 0
        return p

    def _swap(self, col, block):
        if isinstance(block, DictionaryBlock):
            col.ids = self._up(block.ids)
            self._swap(col.dictionary.contents, block.dictionary)
        elif isinstance(block, RunLengthEncodedBlock):
            self._swap(col.dictionary.contents, block.value)
        else:
            col.values = self._up(block.values)
            if block.valueIsNull is not None:
                col.nulls = self._up(block.valueIsNull)

    def free(self):
        for p in self._allocs:
            N.lib().pgpu_device_free(self.device, p)
        self._allocs = []


def concat(pages):
    """one Page out of several flat pages (same block classes)"""
    blocks = []
    for c in range(len(pages[0].blocks)):
        bs = [p.blocks[c] for p in pages]
        nulls = None
        if any(b.valueIsNull is not None for b in bs):
            nulls = np.concatenate([b.valueIsNull if b.valueIsNull is not None else np.zeros(b.position_count, np.uint8) for b in bs])
        blocks.append(type(bs[0])(bs[0].type, np.concatenate([b.values for b in bs]), nulls))
    return Page(*blocks)


def run_swept(plan, page, monkeypatch, expect_sweeps=None, more_pages=(), **knobs):
    monkeypatch.setenv("PGPU_TUNE_SWEEP_MIN_ROWS", str(knobs.pop("min_rows", 1000)))
    monkeypatch.setenv("PGPU_TUNE_SWEEP_MIN_TABLE_MB", str(knobs.pop("min_table_mb", 0)))
    monkeypatch.setenv("PGPU_TUNE_SWEEP_FORCE", str(knobs.pop("force", 1)))   # 0: keys that arrive in runs stay on the ordinary path
    for k, v in knobs.items():
        monkeypatch.setenv("PGPU_TUNE_SWEEP_" + k.upper(), str(v))
    dp = DevicePageOf(page)
    op = GpuPipelineOperator(plan, 0)
    try:
        op.addDeviceInput(dp.cols, dp.ncols, dp.rows)
        out = rows_of(toPages(op, list(more_pages)))
        st = op.stats()
    finally:
        op.close()
        dp.free()
    if expect_sweeps is not None:
        assert st.partition_sweeps == expect_sweeps, st.partition_sweeps
    return out, st


def highcard(shuffled, rows=400_003):
    return tpch.highcard_plan(NO_DIRECTORY), tpch.highcard_page(0, rows, rows, shuffled)


@pytest.mark.gpu
@pytest.mark.parametrize("k", [None, 1, 64, 1024])
@pytest.mark.parametrize("shuffled", [False, True])
def test_sweep_equals_oracle(shuffled, k, monkeypatch):
    plan, page = highcard(shuffled)
    exp = run_oracle(plan, [page])
    knobs = {} if k is None else {"k": k}
    act, st = run_swept(plan, page, monkeypatch, expect_sweeps=1, **knobs)
    assert len(exp) > 90_000
    assert_rows_equal_ignore_order(act, exp, nkeys=1, rtol=0.0)


@pytest.mark.gpu
def test_sweep_only_for_keys_in_no_order(monkeypatch):
    """the clustering probe: l_orderkey as the table holds it (runs of ~4 rows, neighbouring keys) stays on the ordinary path,
    the shuffled column is swept"""
    for shuffled in (False, True):
        plan, page = highcard(shuffled)
        act, st = run_swept(plan, page, monkeypatch, expect_sweeps=1 if shuffled else 0, force=0)
        assert_rows_equal_ignore_order(act, run_oracle(plan, [page]), nkeys=1, rtol=0.0)


@pytest.mark.gpu
@pytest.mark.parametrize("rows", [5, 1023, 4096, 4097, 3 * 4096 + 5])
def test_sweep_ragged_lengths(rows, monkeypatch):
    plan, page = highcard(True, rows)
    act, _ = run_swept(plan, page, monkeypatch, expect_sweeps=1, min_rows=1)
    assert_rows_equal_ignore_order(act, run_oracle(plan, [page]), nkeys=1, rtol=0.0)


def runs_page(n, seed, key_null_frac):
    rng = np.random.default_rng(seed)
    lens = rng.integers(1, 7, n)
    k = np.repeat(rng.integers(-50_000, 50_000, n), lens)[:n]
    knull = np.repeat(rng.random(n) < key_null_frac, lens)[:n]
    k2 = np.repeat(rng.integers(0, 1 << 40, n), rng.integers(1, 4, n))[:n]
    q = rng.integers(1, 51, n).astype(np.float64)
    qnull = rng.random(n) < 0.1
    return Page(IntArrayBlock(INTEGER, k, knull.astype(np.uint8)), LongArrayBlock(BIGINT, k2), LongArrayBlock(DOUBLE, q, qnull.astype(np.uint8)))


@pytest.mark.gpu
@pytest.mark.parametrize("two_keys", [False, True])
def test_sweep_filter_nulls_and_two_key_words(two_keys, monkeypatch):
    """the scatter drops rows the filter rejects (NULL quantity fails it) and moves null flags with the values; NULL keys form
    a group; two key fields = two key words"""
    types = [INTEGER, BIGINT, DOUBLE]
    plan = plan_text(types, compare("lt", field(2, DOUBLE), constant(30.0, DOUBLE)), [field(0, INTEGER), field(1, BIGINT), field(2, DOUBLE)],
                     step="single", groupByChannels=[1, 0] if two_keys else [0],
                     aggregations=[Aggregation("sum", DOUBLE, 2, DOUBLE), Aggregation("count", BIGINT), Aggregation("max", DOUBLE, 2, DOUBLE)],
                     expectedGroups=NO_DIRECTORY)
    page = runs_page(290_003, 7, 0.05)
    act, _ = run_swept(plan, page, monkeypatch, expect_sweeps=1)
    assert_rows_equal_ignore_order(act, run_oracle(plan, [page]), nkeys=2 if two_keys else 1, rtol=0.0)


@pytest.mark.gpu
def test_sweep_no_row_passes_the_filter(monkeypatch):
    plan = plan_text([BIGINT, DOUBLE], compare("lt", field(1, DOUBLE), constant(-1.0, DOUBLE)), [field(0, BIGINT), field(1, DOUBLE)], step="single",
                     groupByChannels=[0], aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE)], expectedGroups=NO_DIRECTORY)
    _, page = highcard(True, 50_000)
    act, _ = run_swept(plan, page, monkeypatch, expect_sweeps=1)
    assert act == []


@pytest.mark.gpu
def test_sweep_dictionary_and_rle_blocks(monkeypatch):
    """dictionary ids are moved, the dictionary stays where it is; an RLE block has nothing to move"""
    rng = np.random.default_rng(3)
    n = 200_001
    dict_keys = LongArrayBlock(BIGINT, rng.permutation(60_000).astype(np.int64) * 7 - 100_000, (rng.random(60_000) < 0.001).astype(np.uint8))
    ids = rng.integers(0, 60_000, n).astype(np.int32)
    q = rng.integers(1, 51, n).astype(np.float64)
    page = Page(DictionaryBlock(dict_keys, ids), LongArrayBlock(DOUBLE, q), RunLengthEncodedBlock(LongArrayBlock(BIGINT, [3]), n))
    plan = plan_text([BIGINT, DOUBLE, BIGINT], None, [field(0, BIGINT), field(1, DOUBLE), field(2, BIGINT)], step="single", groupByChannels=[0],
                     aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE), Aggregation("sum", BIGINT, 2, BIGINT)], expectedGroups=NO_DIRECTORY)
    act, _ = run_swept(plan, page, monkeypatch, expect_sweeps=1)
    assert_rows_equal_ignore_order(act, run_oracle(plan, [page]), nkeys=1, rtol=0.0)


@pytest.mark.gpu
def test_sweep_long_decimal_values(monkeypatch):
    """16-byte values move as two words; decimal sums are exact"""
    rng = np.random.default_rng(5)
    n = 150_000
    d30 = createDecimalType(30, 2)
    k = rng.integers(0, 40_000, n).astype(np.int64)
    vals = [int(x) * 10**15 + int(y) for x, y in zip(rng.integers(-10**9, 10**9, n), rng.integers(0, 10**6, n))]
    page = Page(LongArrayBlock(BIGINT, k), Int128ArrayBlock(d30, vals, (rng.random(n) < 0.05).astype(np.uint8)))
    plan = plan_text([BIGINT, d30], step="single", groupByChannels=[0], aggregations=[Aggregation("sum", createDecimalType(38, 2), 1, d30), Aggregation("count", BIGINT, 1, d30)],
                     expectedGroups=NO_DIRECTORY)
    act, _ = run_swept(plan, page, monkeypatch, expect_sweeps=1)
    assert_rows_equal_ignore_order(act, run_oracle(plan, [page]), nkeys=1, rtol=0.0)


@pytest.mark.gpu
def test_sweep_table_growth_and_batches(monkeypatch):
    """1.7 M distinct keys into a table sized for 1.05 M, in batches of 700 000 rows: rows past the load-factor limit are deferred
    while a batch is swept, the table is grown (its slices move, the partition order of later batches follows the new capacity)
    and the rows are replayed"""
    rng = np.random.default_rng(11)
    distinct = 1_700_000
    keys = np.concatenate([rng.permutation(distinct), np.repeat(rng.permutation(distinct)[:300_000], 2)]).astype(np.int64) * 3 + 1
    q = rng.integers(1, 51, len(keys)).astype(np.float64)
    page = Page(LongArrayBlock(BIGINT, keys), LongArrayBlock(DOUBLE, q))
    monkeypatch.setenv("PGPU_TUNE_SWEEP_MIN_ROWS", "1000")
    monkeypatch.setenv("PGPU_TUNE_SWEEP_MIN_TABLE_MB", "0")
    monkeypatch.setenv("PGPU_TUNE_SWEEP_BATCH_ROWS", "700000")
    dp = DevicePageOf(page)
    op = GpuPipelineOperator(tpch.highcard_plan(NO_DIRECTORY), 0)
    try:
        op.addDeviceInput(dp.cols, 2, dp.rows)
        out = toPages(op, [])
        st = op.stats()
    finally:
        op.close()
        dp.free()
    assert st.partition_sweeps == 4 and st.table_rehashes >= 1
    k = np.concatenate([p.blocks[0].values for p in out])
    s = np.concatenate([p.blocks[1].values.view(np.float64) for p in out])
    order = np.argsort(k, kind="stable")
    uk, inv = np.unique(keys, return_inverse=True)
    assert np.array_equal(k[order], uk)
    assert np.array_equal(s[order], np.bincount(inv, weights=q))


@pytest.mark.gpu
def test_sweep_mixes_with_ordinary_pages_and_survives_reset(monkeypatch):
    plan, page = highcard(True, 300_000)
    more = [tpch.highcard_page(300_000, 50_000, 350_000, True), tpch.highcard_page(0, 20_000, 350_000, False)]
    exp = run_oracle(plan, [page] + more)
    act, _ = run_swept(plan, page, monkeypatch, expect_sweeps=1, more_pages=more)
    assert_rows_equal_ignore_order(act, exp, nkeys=1, rtol=0.0)
    # the same pipeline again after reset: the reordered copy and the table are reused
    dp = DevicePageOf(page)
    op = GpuPipelineOperator(plan, 0)
    try:
        for _ in range(2):
            op.addDeviceInput(dp.cols, 2, dp.rows)
            again = rows_of(toPages(op, more))
            assert_rows_equal_ignore_order(again, exp, nkeys=1, rtol=0.0)
            assert op.stats().partition_sweeps == 1
            op.reset()
    finally:
        op.close()
        dp.free()


@pytest.mark.gpu
def test_sweep_thresholds_and_switch(monkeypatch):
    """defaults: a 400 000-row page and a 50 MB table stay on the ordinary path; PGPU_TUNE_SWEEP=0 removes the kernels"""
    plan, page = highcard(True)
    exp = run_oracle(plan, [page])
    dp = DevicePageOf(page)
    try:
        for env, sweeps in (({}, 0), ({"PGPU_TUNE_SWEEP_MIN_ROWS": "1000", "PGPU_TUNE_SWEEP_MIN_TABLE_MB": "0"}, 1),
                            ({"PGPU_TUNE_SWEEP_MIN_ROWS": "1000", "PGPU_TUNE_SWEEP_MIN_TABLE_MB": "0", "PGPU_TUNE_SWEEP": "0"}, 0)):
            with monkeypatch.context() as m:
                for k in ("PGPU_TUNE_SWEEP", "PGPU_TUNE_SWEEP_MIN_ROWS", "PGPU_TUNE_SWEEP_MIN_TABLE_MB"):
                    m.delenv(k, raising=False)
                for k, v in env.items():
                    m.setenv(k, v)
                op = GpuPipelineOperator(plan, 0)
                try:
                    op.addDeviceInput(dp.cols, 2, dp.rows)
                    act = rows_of(toPages(op, []))
                    assert op.stats().partition_sweeps == sweeps
                    assert ("pgpu_k_page_sweep" in op.generatedSource()) == (env.get("PGPU_TUNE_SWEEP") != "0")
                finally:
                    op.close()
                assert_rows_equal_ignore_order(act, exp, nkeys=1, rtol=0.0)
    finally:
        dp.free()


def test_sweep_kernels_compile_without_gpu():
    for plan, sig in ((tpch.highcard_plan(NO_DIRECTORY), ([0, 0], [False, False])),
                      (plan_text([INTEGER, BIGINT, DOUBLE], compare("lt", field(2, DOUBLE), constant(30.0, DOUBLE)), [field(0, INTEGER), field(1, BIGINT), field(2, DOUBLE)],
                                 step="single", groupByChannels=[1, 0], aggregations=[Aggregation("sum", DOUBLE, 2, DOUBLE)], expectedGroups=NO_DIRECTORY),
                       ([0, 0, 0], [True, False, True]))):
        op = GpuPipelineOperator(plan, -1)
        try:
            op.compileSignature(*sig)
            src = op.generatedSource()
            assert "pgpu_k_part_count" in src and "pgpu_k_part_scatter" in src and "pgpu_k_page_sweep" in src and "move_tile" in src
        finally:
            op.close()
    # a var-width (offsets + bytes) key column cannot be moved row by row: no sweep kernels, the ordinary path stays
    from presto_b200.spi import createVarcharType
    v5 = createVarcharType(5)
    op = GpuPipelineOperator(plan_text([v5, DOUBLE], step="single", groupByChannels=[0], aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE)], expectedGroups=NO_DIRECTORY), -1)
    try:
        op.compileSignature([0, 0], [False, False])
        assert "pgpu_k_page_sweep" not in op.generatedSource()
    finally:
        op.close()
