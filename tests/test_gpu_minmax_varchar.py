"""min / max over VARCHAR(n <= 15) (AbstractMinMaxAggregationFunction on Slice values: MaxAggregationFunction / SliceState,
compared with Slice.compareTo = unsigned bytes, a proper prefix first): the value's 16-byte cell mapped to an order-preserving
128-bit image in the same compare-and-swap cell as long DECIMAL min / max (csrc/device/pgpu_device.cuh varchar_order_image).
Global, shared-memory directory, HBM table, PARTIAL -> FINAL, dictionary blocks; against the oracle and Python's bytes order."""
import numpy as np
import pytest

from presto_b200.operators import Aggregation, GpuPipelineOperator, plan_text, toPages
from presto_b200.spi import BIGINT, DictionaryBlock, LongArrayBlock, Page, VariableWidthBlock, createVarcharType
from util import assert_rows_equal_ignore_order, rows_of, run_gpu, run_oracle

pytestmark = pytest.mark.gpu
V15, V3 = createVarcharType(15), createVarcharType(3)
NO_DIRECTORY = (1 << 20) + 1
ALPHABET = ["", "a", "A", "b", "ab", "a ", "\x01", "\x7f", "é", "ÿ", "zz", "z"]   # multi-byte UTF-8 orders by its bytes


def strings(n, seed, t):
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        parts, size = [], 0
        for _ in range(int(rng.integers(0, 14))):
            w = ALPHABET[int(rng.integers(0, len(ALPHABET)))]
            if size + len(w.encode()) > t.precision or len("".join(parts)) + len(w) > t.precision:
                break
            parts.append(w)
            size += len(w.encode())
        out.append(None if rng.random() < 0.1 else "".join(parts))
    return out


def exact(keys, vals):
    groups = {}
    for k, v in zip(keys, vals):
        lo, hi, cnt = groups.get(k, (None, None, 0))
        if v is not None:
            b = v.encode()
            lo = v if lo is None or b < lo.encode() else lo
            hi = v if hi is None or b > hi.encode() else hi
        groups[k] = (lo, hi, cnt + 1)
    return sorted(((k, lo, hi, c) for k, (lo, hi, c) in groups.items()), key=lambda r: r[0])


def plan_for(t, expected_groups, step="single"):
    return plan_text([BIGINT, t], step=step, groupByChannels=[0],
                     aggregations=[Aggregation("min", t, 1, t), Aggregation("max", t, 1, t), Aggregation("count", BIGINT)], expectedGroups=expected_groups)


@pytest.mark.parametrize("t", [V15, V3], ids=["varchar15", "varchar3"])
def test_global_min_max_varchar(t):
    vals = strings(30_001, 3, t)
    page = Page(VariableWidthBlock.from_values(t, vals))
    plan = plan_text([t], step="single", aggregations=[Aggregation("min", t, 0, t), Aggregation("max", t, 0, t)])
    live = [v for v in vals if v is not None]
    want = [(min(live, key=str.encode), max(live, key=str.encode))]
    assert run_gpu(plan, [page]) == want == run_oracle(plan, [page])
    empty = Page(VariableWidthBlock.from_values(t, [None, None]))
    assert run_gpu(plan, [empty]) == [(None, None)] == run_oracle(plan, [empty])


@pytest.mark.parametrize("expected_groups,ngroups", [(100, 37), (NO_DIRECTORY, 20_000)], ids=["directory", "hbm_table"])
def test_grouped_min_max_varchar(expected_groups, ngroups):
    rng = np.random.default_rng(5)
    n = 150_003
    keys = rng.integers(0, ngroups, n)
    vals = strings(n, 7, V15)
    for i in np.flatnonzero(keys == 3):
        vals[i] = None                       # one group whose values are all NULL
    page = Page(LongArrayBlock(BIGINT, keys), VariableWidthBlock.from_values(V15, vals))
    plan = plan_for(V15, expected_groups)
    got = run_gpu(plan, [page])
    assert sorted(got, key=lambda r: r[0]) == exact(keys.tolist(), vals)
    assert_rows_equal_ignore_order(got, run_oracle(plan, [page]), nkeys=1, rtol=0.0)


def test_min_max_varchar_dictionary_block_and_partial_final():
    rng = np.random.default_rng(9)
    words = sorted(set(s for s in strings(400, 11, V15) if s is not None))
    n = 80_000
    keys = rng.integers(0, 3_000, n)
    ids = rng.integers(0, len(words), n).astype(np.int32)
    vals = [words[i] for i in ids]
    pages = [Page(LongArrayBlock(BIGINT, keys[:n // 2]), DictionaryBlock(VariableWidthBlock.from_values(V15, words), ids[:n // 2])),
             Page(LongArrayBlock(BIGINT, keys[n // 2:]), VariableWidthBlock.from_values(V15, vals[n // 2:]))]
    mids = []
    for pg in pages:
        op = GpuPipelineOperator(plan_for(V15, NO_DIRECTORY, step="partial"), 0)
        try:
            mids += toPages(op, [pg])
        finally:
            op.close()
    final = plan_text([BIGINT, V15, V15, BIGINT], step="final", groupByChannels=[0],
                      aggregations=[Aggregation("min", V15, 1, V15), Aggregation("max", V15, 2, V15), Aggregation("count", BIGINT, 3, BIGINT)], expectedGroups=NO_DIRECTORY)
    fop = GpuPipelineOperator(final, 0)
    try:
        got = rows_of(toPages(fop, mids))
    finally:
        fop.close()
    assert sorted(got, key=lambda r: r[0]) == exact(keys.tolist(), vals)
    single = plan_for(V15, NO_DIRECTORY)
    assert_rows_equal_ignore_order(run_gpu(single, pages), run_oracle(single, pages), nkeys=1, rtol=0.0)
