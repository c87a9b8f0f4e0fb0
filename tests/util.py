"""Shared helpers of the test-suite: drive an operator by hand (OperatorAssertion.toPages of the reference) and
compare result pages ignoring row order (assertPagesEqualIgnoreOrder, TestHashAggregationOperator.java:216)."""
import math

from presto_b200.operators import GpuPipelineOperator, toPages
from oracle_op import OracleOperator

DOUBLE_RTOL = 1e-12  # north_star: DOUBLE sums within 1e-12 relative (summation order differs)


def rows_of(pages):
    out = []
    for p in pages:
        out.extend(p.to_rows())
    return out


def run_oracle(plan, pages):
    op = OracleOperator(plan)
    try:
        return rows_of(toPages(op, pages))
    finally:
        op.close()


def run_gpu(plan, pages, device=0):
    op = GpuPipelineOperator(plan, device)
    try:
        return rows_of(toPages(op, pages))
    finally:
        op.close()


def _key(row, nkeys):
    return tuple((v is None, 0 if v is None else v) for v in row[:nkeys])


def _sort_key(row):
    out = []
    for v in row:
        if v is None:
            out.append((2, 0))
        elif isinstance(v, float):
            out.append((1, 0.0 if math.isnan(v) else round(v, 3)))
        elif isinstance(v, (bytes, str)):
            out.append((1, v if isinstance(v, str) else v.decode("latin1")))
        else:
            out.append((1, v))
    return tuple(str(x) for x in out)


def value_equal(a, b, rtol=DOUBLE_RTOL):
    if a is None or b is None:
        return a is None and b is None
    if isinstance(a, float) or isinstance(b, float):
        a, b = float(a), float(b)
        if math.isnan(a) or math.isnan(b):
            return math.isnan(a) and math.isnan(b)
        if a == b:
            return True
        return abs(a - b) <= rtol * max(abs(a), abs(b))
    return a == b


def assert_rows_equal_ignore_order(actual, expected, nkeys=None, rtol=DOUBLE_RTOL):
    """nkeys: leading grouping columns used to pair rows (exact match); None -> sort whole rows."""
    assert len(actual) == len(expected), "row count %d != %d" % (len(actual), len(expected))
    if nkeys:
        am = {}
        for r in actual:
            k = _key(r, nkeys)
            assert k not in am, "duplicate group %r" % (k,)
            am[k] = r
        for e in expected:
            k = _key(e, nkeys)
            assert k in am, "missing group %r" % (e[:nkeys],)
            a = am[k]
            for j, (x, y) in enumerate(zip(a, e)):
                assert value_equal(x, y, rtol), "group %r column %d: %r != %r" % (e[:nkeys], j, x, y)
    else:
        sa, se = sorted(actual, key=_sort_key), sorted(expected, key=_sort_key)
        for a, e in zip(sa, se):
            assert len(a) == len(e)
            for j, (x, y) in enumerate(zip(a, e)):
                assert value_equal(x, y, rtol), "row %r vs %r column %d" % (a, e, j)


def assert_rows_equal_ordered(actual, expected, rtol=0.0):
    assert len(actual) == len(expected), "row count %d != %d" % (len(actual), len(expected))
    for i, (a, e) in enumerate(zip(actual, expected)):
        for j, (x, y) in enumerate(zip(a, e)):
            assert value_equal(x, y, rtol), "row %d column %d: %r != %r" % (i, j, x, y)


# ---------------------------------------------------------------------------------------------- exact references
# DOUBLE sums depend on the summation order: the reference adds sequentially in position order (error grows like
# n * 2^-53), the GPU adds in a tree. Both are therefore compared with the correctly rounded sum (math.fsum) of
# the same per-row values; the per-row projections are plain IEEE operations and are bit-identical everywhere.
def sequential_bound(n):
    """relative error bound of a sequential double sum of n same-signed terms"""
    return max(DOUBLE_RTOL, n * 2.0 ** -53)


def exact_q1(pages):
    import numpy as np
    groups = {}
    for p in pages:
        rf, ls = p.blocks[0].to_pylist(), p.blocks[1].to_pylist()
        col = lambda c: np.array(p.blocks[c].to_pylist(), dtype=np.float64)
        q, e, d, t = col(2), col(3), col(4), col(5)
        ship = np.array(p.blocks[6].to_pylist())
        dp = e * (1.0 - d)
        ch = dp * (1.0 + t)
        sel = ship <= 10471
        keys = np.array([a + b for a, b in zip(rf, ls)])
        for k in np.unique(keys[sel]):
            m = sel & (keys == k)
            g = groups.setdefault(k, [[], [], [], [], [], 0])
            for i, v in enumerate((q, e, dp, ch, d)):
                g[i].append(v[m])
            g[5] += int(m.sum())
    rows = []
    for k, g in groups.items():
        s = [math.fsum(np.concatenate(x).tolist()) for x in g[:5]]
        n = g[5]
        rows.append((k[0], k[1], s[0], s[1], s[2], s[3], s[0] / n, s[1] / n, s[4] / n, n))
    return rows


def exact_q6(pages):
    import numpy as np
    parts = []
    for p in pages:
        col = lambda c: np.array(p.blocks[c].to_pylist(), dtype=np.float64)
        e, d, q = col(0), col(1), col(3)
        ship = np.array(p.blocks[2].to_pylist())
        sel = (ship >= 8766) & (ship < 9131) & (d >= 0.05) & (d <= 0.07) & (q < 24.0)
        parts.append((e * d)[sel])
    v = np.concatenate(parts)
    return [(math.fsum(v.tolist()) if len(v) else None,)], len(v)
