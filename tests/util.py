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
