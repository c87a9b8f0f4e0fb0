"""Known-answer cases restated from the reference's own tests (paths under /root/reference/core/trino-main/src/test/java/io/trino/):

  aggregate KATs   operator/aggregation/AbstractTestAggregationFunction.java:84-143 — every aggregate is run over the
                   sequence blocks start..start+length for (0,0) (0,1) (0,5) (-10,5) (2,4), over 10 NULLs, and over 10
                   values alternating with NULLs; expectations are the closed forms of Test{DoubleSum,LongSum,Count,
                   CountColumn,DoubleAverage,LongAverage,LongMin,LongMax,DoubleMin,DoubleMax,LongDecimalMin,LongDecimalMax,ShortDecimalSum,LongDecimalSum,
                   ShortDecimalAverage,LongDecimalAverage}Aggregation.java#getExpectedValue
  hash aggregation operator/TestHashAggregationOperator.java:161-219 (40 000 groups x 3 pages -> (i, 3, 3i, i, 3, 3)),
                   :222-271 (global default row over empty input), :477-560 (multi-page output)
  page processor   operator/project/TestPageProcessor.java:125-200, TestDictionaryAwarePageFilter.java:49-116
  NULLIF / CASE    sql/gen/TestExpressionCompiler.java:321-348,440,492 (nullif over the left x right grids, expectation
                   #nullIf :675-697), :981-1024 (simple CASE over value x first x second grids), :1649-1655;
                   operator/scalar/TestConditions.java:308-380 (simple CASE literals, NULL operands and NULL value)

Each case = (name, plan text, input pages, expected rows, number of leading key columns or None).
The same cases run against the CPU oracle (-m "not gpu": pins the oracle) and the CUDA path (-m gpu: parity).
"""
import math

import numpy as np

from presto_b200.operators import Aggregation, plan_text
from presto_b200.relational import and_, call, compare, constant, field, or_, SpecialForm
from presto_b200.spi import (BIGINT, BOOLEAN, DATE, DOUBLE, INTEGER, DictionaryBlock, IntArrayBlock, LongArrayBlock, Page,
                             RunLengthEncodedBlock, VariableWidthBlock, block_of, createDecimalType, createVarcharType)

SEQUENCES = [(0, 0), (0, 1), (0, 5), (-10, 5), (2, 4)]


def _seq(start, length):
    return list(range(start, start + length))


def _avg_decimal(values, scale_div=1):
    """BigDecimal sum / count, ROUND_HALF_UP at the input scale (DecimalAverageAggregation.java:218-227)."""
    if not values:
        return None
    s, n = sum(values), len(values)
    q, r = divmod(abs(s), n)
    if 2 * r >= n:
        q += 1
    return -q if s < 0 else q


def aggregate_function_cases():
    """(name, type, to_block_values, aggregations, expected(values) -> row)"""
    dec_short = createDecimalType(10, 2)   # TestShortDecimalSumAggregation: DECIMAL(10, 2)... values i as unscaled
    dec_long = createDecimalType(30, 5)
    dec_sum_out = lambda t: createDecimalType(38, t.scale)
    fns = []

    def add(name, t, conv, aggs, expect):
        fns.append((name, t, conv, aggs, expect))

    add("sum_double", DOUBLE, float, [Aggregation("sum", DOUBLE, 0, DOUBLE)], lambda v: (math.fsum(v) if v else None,))
    add("sum_bigint", BIGINT, int, [Aggregation("sum", BIGINT, 0, BIGINT)], lambda v: (sum(v) if v else None,))
    add("count_star", BIGINT, int, [Aggregation("count", BIGINT)], None)  # expectation needs the row count: special-cased
    add("count_column", BIGINT, int, [Aggregation("count", BIGINT, 0, BIGINT)], lambda v: (len(v),))
    add("avg_double", DOUBLE, float, [Aggregation("avg", DOUBLE, 0, DOUBLE)], lambda v: (math.fsum(v) / len(v) if v else None,))
    add("avg_bigint", BIGINT, int, [Aggregation("avg", DOUBLE, 0, BIGINT)], lambda v: (sum(v) / len(v) if v else None,))
    add("min_bigint", BIGINT, int, [Aggregation("min", BIGINT, 0, BIGINT)], lambda v: (min(v) if v else None,))
    add("max_bigint", BIGINT, int, [Aggregation("max", BIGINT, 0, BIGINT)], lambda v: (max(v) if v else None,))
    add("min_double", DOUBLE, float, [Aggregation("min", DOUBLE, 0, DOUBLE)], lambda v: (min(v) if v else None,))
    add("max_double", DOUBLE, float, [Aggregation("max", DOUBLE, 0, DOUBLE)], lambda v: (max(v) if v else None,))
    add("min_integer", INTEGER, int, [Aggregation("min", INTEGER, 0, INTEGER)], lambda v: (min(v) if v else None,))
    add("max_date", DATE, int, [Aggregation("max", DATE, 0, DATE)], lambda v: (max(v) if v else None,))
    # TestLongDecimalMinAggregation / TestLongDecimalMaxAggregation: DECIMAL(30, 5), BigDecimal.valueOf(i) = unscaled i * 10^5;
    # expected start / start + length - 1. Large magnitudes of both signs are in tests/test_gpu_operators.py.
    big = lambda x: int(x) * 10**5
    add("min_long_decimal", dec_long, big, [Aggregation("min", dec_long, 0, dec_long)], lambda v: (min(v) if v else None,))
    add("max_long_decimal", dec_long, big, [Aggregation("max", dec_long, 0, dec_long)], lambda v: (max(v) if v else None,))
    add("min_short_decimal", dec_short, int, [Aggregation("min", dec_short, 0, dec_short)], lambda v: (min(v) if v else None,))
    add("max_short_decimal", dec_short, int, [Aggregation("max", dec_short, 0, dec_short)], lambda v: (max(v) if v else None,))
    add("sum_short_decimal", dec_short, int, [Aggregation("sum", dec_sum_out(dec_short), 0, dec_short)], lambda v: (sum(v) if v else None,))
    add("sum_long_decimal", dec_long, int, [Aggregation("sum", dec_sum_out(dec_long), 0, dec_long)], lambda v: (sum(v) if v else None,))
    add("avg_short_decimal", dec_short, int, [Aggregation("avg", dec_short, 0, dec_short)], lambda v: (_avg_decimal(v),))
    add("avg_long_decimal", dec_long, int, [Aggregation("avg", dec_long, 0, dec_long)], lambda v: (_avg_decimal(v),))
    return fns


def global_aggregate_cases():
    cases = []
    for name, t, conv, aggs, expect in aggregate_function_cases():
        plan = plan_text([t], step="single", aggregations=aggs)
        variants = [("seq_%d_%d" % s, [conv(x) for x in _seq(*s)]) for s in SEQUENCES]
        variants.append(("all_null", [None] * 10))
        alt = []
        for x in _seq(0, 10):
            alt += [None, conv(x)]
        variants.append(("alternating_nulls", alt))
        for vname, vals in variants:
            nn = [x for x in vals if x is not None]
            exp = (len(vals),) if expect is None else expect(nn)
            pages = [Page(block_of(t, vals))] if vals else [Page(block_of(t, []), positionCount=0)]
            cases.append(("%s/%s" % (name, vname), plan, pages, [exp], None))
        # the same values split over two pages (AbstractTestAggregationFunction.testSlidingWindow spirit: state carries over)
        vals = [conv(x) for x in _seq(-3, 9)]
        exp = (len(vals),) if expect is None else expect(vals)
        cases.append(("%s/two_pages" % name, plan, [Page(block_of(t, vals[:4])), Page(block_of(t, vals[4:]))], [exp], None))
    return cases


def hash_aggregation_cases(number_of_rows=40_000):
    """TestHashAggregationOperator.testHashAggregation (:161-219): three sequence pages of `number_of_rows` rows whose key
    channel is the row index; expected per group i: count 3, sum 3i, avg i, count(col) 3, count(bool) 3.
    (The reference declares its VARCHAR channels unbounded; the GPU subset holds varchar(n <= 15) values, so the key and the
    max(varchar) channel are varchar(8) of the same digits: same rows, same expected strings.)"""
    V8 = createVarcharType(8)
    n = number_of_rows
    keys = [str(i) for i in range(n)]
    pages = []
    for base in (100_000, 200_000, 300_000):
        pages.append(Page(VariableWidthBlock.from_values(V8, keys), LongArrayBlock(BIGINT, np.arange(n)),
                          VariableWidthBlock.from_values(V8, [str(base + i) for i in range(n)]),
                          block_of(BOOLEAN, [(i + 500) % 2 == 0 for i in range(n)])))
    aggs = [Aggregation("count", BIGINT), Aggregation("sum", BIGINT, 1, BIGINT), Aggregation("avg", DOUBLE, 1, BIGINT),
            Aggregation("max", V8, 2, V8), Aggregation("count", BIGINT, 0, V8), Aggregation("count", BIGINT, 3, BOOLEAN)]
    plan = plan_text([V8, BIGINT, V8, BOOLEAN], step="single", groupByChannels=[0], aggregations=aggs, expectedGroups=100_000,
                     maxOutputRows=8192)
    exp = [(str(i), 3, 3 * i, float(i), str(300_000 + i), 3, 3) for i in range(n)]
    cases = [("hash_aggregation_40000x3", plan, pages, exp, 1)]

    # BIGINT key with a NULL group (TestGroupByHash.testNullGroup :157-185) + rehash from a tiny expected size (:225-250)
    kv = [None if i % 7 == 0 else i % 1000 for i in range(20_000)]
    vals = [float(i) for i in range(20_000)]
    plan = plan_text([BIGINT, DOUBLE], step="single", groupByChannels=[0],
                     aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE), Aggregation("count", BIGINT)], expectedGroups=4)
    groups = {}
    for k, v in zip(kv, vals):
        g = groups.setdefault(k, [[], 0])
        g[0].append(v)
        g[1] += 1
    exp = [(k, math.fsum(g[0]), g[1]) for k, g in groups.items()]
    cases.append(("bigint_key_null_group_rehash", plan, [Page(block_of(BIGINT, kv[:5000]), block_of(DOUBLE, vals[:5000])),
                                                        Page(block_of(BIGINT, kv[5000:]), block_of(DOUBLE, vals[5000:]))], exp, 1))

    # multi-channel key (MultiChannelGroupByHash): (varchar(1), date, double) incl. -0.0/+0.0 and NaN keys, which group by
    # IS NOT DISTINCT FROM (spi/type/DoubleType.java:163-195)
    V1 = createVarcharType(1)
    a = ["A", "B", None, "A", "B", None, "A", "A"] * 50
    d = [1, 2, 3, 1, 2, 3, 1, 9] * 50
    f = [0.0, float("nan"), 1.5, -0.0, float("nan"), 1.5, 0.0, 2.0] * 50
    x = list(range(400))
    plan = plan_text([V1, DATE, DOUBLE, BIGINT], step="single", groupByChannels=[0, 1, 2],
                     aggregations=[Aggregation("sum", BIGINT, 3, BIGINT), Aggregation("count", BIGINT)], expectedGroups=16)
    groups = {}
    for ka, kd, kf, v in zip(a, d, f, x):
        kk = (ka, kd, "nan" if kf != kf else (0.0 if kf == 0 else kf))
        g = groups.setdefault(kk, [ka, kd, kf, 0, 0])
        g[3] += v
        g[4] += 1
    exp = [tuple(g) for g in groups.values()]
    cases.append(("multi_channel_key_nan_zero", plan, [Page(block_of(V1, a), block_of(DATE, d), block_of(DOUBLE, f), block_of(BIGINT, x))], exp, None))

    # global aggregation over no input still yields one default row (TestHashAggregationOperator :222-271 /
    # TestAggregationOperator): count 0, sums NULL
    plan = plan_text([BIGINT, DOUBLE], step="single", aggregations=[Aggregation("count", BIGINT), Aggregation("sum", BIGINT, 0, BIGINT),
                                                                     Aggregation("avg", DOUBLE, 1, DOUBLE), Aggregation("max", DOUBLE, 1, DOUBLE)])
    cases.append(("global_default_row", plan, [], [(0, None, None, None)], None))
    # grouped aggregation over no input yields no rows (HashAggregationOperator.java:449-463)
    plan = plan_text([BIGINT, DOUBLE], step="single", groupByChannels=[0], aggregations=[Aggregation("count", BIGINT)])
    cases.append(("grouped_no_input", plan, [], [], None))

    # masked aggregates (AccumulatorCompiler.java:535-553): count(*) FILTER (mask), sum(x) FILTER (mask)
    m = [True, False, None, True, True, False] * 20
    v = list(range(120))
    k = [i % 3 for i in range(120)]
    plan = plan_text([BIGINT, BIGINT, BOOLEAN], step="single", groupByChannels=[0],
                     aggregations=[Aggregation("count", BIGINT, None, None, 2), Aggregation("sum", BIGINT, 1, BIGINT, 2), Aggregation("count", BIGINT)])
    exp = []
    for g in range(3):
        rows = [(vv, mm) for kk, vv, mm in zip(k, v, m) if kk == g]
        sel = [vv for vv, mm in rows if mm]
        exp.append((g, len(sel), sum(sel) if sel else None, len(rows)))
    cases.append(("masked_aggregates", plan, [Page(block_of(BIGINT, k), block_of(BIGINT, v), block_of(BOOLEAN, m))], exp, 1))
    return cases


def _sel(rows, pred):
    return [r for r in rows if pred(r) is True]


def filter_project_cases():
    """TestPageProcessor (:125-200): filter -> SelectedPositions (list and range), projections evaluated at the selected
    positions only, NULL filter result drops the row (PageFunctionCompiler.java:542)."""
    cases = []
    n = 3000
    a = [None if i % 11 == 0 else i - 1500 for i in range(n)]
    b = [None if i % 13 == 0 else (i % 97) * 0.25 for i in range(n)]
    page = Page(block_of(BIGINT, a), block_of(DOUBLE, b))
    A, B = field(0, BIGINT), field(1, DOUBLE)
    # filter: a > -100 AND (b < 10.0 OR a IS NULL)  — three-valued logic
    f = and_(compare("gt", A, constant(-100, BIGINT)), or_(compare("lt", B, constant(10.0, DOUBLE)), SpecialForm("IS_NULL", BOOLEAN, [A])))
    proj = [A, call("multiply", DOUBLE, B, constant(2.0, DOUBLE)), call("add", BIGINT, A, constant(7, BIGINT)),
            SpecialForm("COALESCE", DOUBLE, [B, constant(-1.0, DOUBLE)]),
            SpecialForm("IF", BIGINT, [compare("ge", A, constant(0, BIGINT)), A, call("negate", BIGINT, A)])]
    plan = plan_text([BIGINT, DOUBLE], f, proj)

    def tv_and(x, y):
        if x is False or y is False:
            return False
        if x is None or y is None:
            return None
        return True

    def tv_or(x, y):
        if x is True or y is True:
            return True
        if x is None or y is None:
            return None
        return False

    exp = []
    for x, y in zip(a, b):
        c1 = None if x is None else x > -100
        c2 = tv_or(None if y is None else y < 10.0, x is None)
        if tv_and(c1, c2) is True:
            exp.append((x, None if y is None else y * 2.0, None if x is None else x + 7, -1.0 if y is None else y,
                        None if x is None else (x if x >= 0 else -x)))
    cases.append(("filter_project_three_valued", plan, [page], exp, "ordered"))
    # all selected (range) / none selected
    plan = plan_text([BIGINT, DOUBLE], compare("ge", SpecialForm("COALESCE", BIGINT, [A, constant(0, BIGINT)]), constant(-5000, BIGINT)), [A, B])
    cases.append(("filter_selects_all", plan, [page], list(zip(a, b)), "ordered"))
    plan = plan_text([BIGINT, DOUBLE], compare("gt", A, constant(10 ** 9, BIGINT)), [A, B])
    cases.append(("filter_selects_none", plan, [page], [], "ordered"))
    # BETWEEN and IN on DATE, projection-only pipeline without a filter
    d = [8000 + (i * 37) % 3000 for i in range(n)]
    D = field(0, DATE)
    f = and_(SpecialForm("BETWEEN", BOOLEAN, [D, constant(8766, DATE), constant(9130, DATE)]),
             SpecialForm("IN", BOOLEAN, [field(1, INTEGER), constant(1, INTEGER), constant(3, INTEGER), constant(5, INTEGER)]))
    q = [i % 7 for i in range(n)]
    plan = plan_text([DATE, INTEGER], f, [D, field(1, INTEGER)])
    exp = [(x, y) for x, y in zip(d, q) if 8766 <= x <= 9130 and y in (1, 3, 5)]
    cases.append(("between_in", plan, [Page(block_of(DATE, d), block_of(INTEGER, q))], exp, "ordered"))
    return cases


def conditional_cases():
    """NULL_IF (NullIfCodeGenerator.java:70-113) and SWITCH/WHEN (SwitchCodeGenerator.java:113-159) over the value grids of
    TestExpressionCompiler (:114-136). Every grid combination is one row; the CASE results 'first' / 'second' / 'else' of the
    reference test are the BIGINT codes 1 / 2 / 0 here (VARCHAR values are not computed on this path)."""
    cases = []

    def null_if(left, right):  # TestExpressionCompiler.java:675-697
        if left is None:
            return None
        if right is None:
            return left
        return None if left == right else left

    def grid(*axes):
        rows = [()]
        for ax in axes:
            rows = [r + (v,) for r in rows for v in ax]
        return [list(c) for c in zip(*rows)]

    small_ints = [9, 10, 11, -9, -10, -11, None]
    int_lefts, int_rights, int_middle = small_ints + [101510, 2 ** 31 - 1], [3, -3, 101510823, None], [9, -3, 88, None]
    long_lefts, long_rights = [9, 10, 11, -9, -10, -11, None], [3, -3, 10151082135029369, None]
    dbl_lefts = [9.0, 10.0, 11.0, -9.0, -10.0, -11.0, 9.1, 10.1, 11.1, -9.1, -10.1, -11.1, 5e-324, 1.7976931348623157e308, 2.2250738585072014e-308, None]
    dbl_rights, dbl_middle = [3.0, -3.0, 3.1, -3.1, None], [9.0, -3.1, 88.0, None]
    D31 = createDecimalType(3, 1)
    dec_lefts, dec_rights = [90, 100, 110, -90, -100, -110, 91, 101, 111, -91, -101, -111, None], [30, -30, 31, -31, None]  # unscaled

    # nullif(left, right) for every pair of a grid, and against the left column itself (always NULL or left IS NULL)
    for name, t, lefts, rights in (("integer", INTEGER, int_lefts, int_rights + [9, -11]), ("bigint", BIGINT, long_lefts, long_rights + [10, -9]),
                                   ("double", DOUBLE, dbl_lefts, dbl_rights + [9.1, -0.0, 5e-324]), ("decimal", D31, dec_lefts, dec_rights + [91, -110])):
        lcol, rcol = grid(lefts, rights)
        L, R = field(0, t), field(1, t)
        plan = plan_text([t, t], None, [SpecialForm("NULL_IF", t, [L, R]), SpecialForm("NULL_IF", t, [L, L]),
                                        SpecialForm("NULL_IF", t, [L, constant(None, t)]), SpecialForm("NULL_IF", t, [constant(None, t), R])])
        exp = [(null_if(x, y), None, x, None) for x, y in zip(lcol, rcol)]
        cases.append(("nullif_grid_" + name, plan, [Page(block_of(t, lcol), block_of(t, rcol))], exp, "ordered"))

    # case value when first then 1 when second then 2 else 0 end (doubles); without ELSE (integers, decimals)
    def simple_case(value, first, second, otherwise):
        if value is None:
            return otherwise
        if first is not None and value == first:
            return 1
        if second is not None and value == second:
            return 2
        return otherwise

    for name, t, lefts, middle, rights, with_else in (("double_else", DOUBLE, dbl_lefts, dbl_middle + [11.1], dbl_rights + [-9.0], True),
                                                      ("integer_no_else", INTEGER, int_lefts, int_middle, int_rights + [9], False),
                                                      ("decimal_no_else", D31, dec_lefts, [90, -31, 880, None], dec_rights + [90], False)):
        v, f, s2 = grid(lefts, middle, rights)
        args = [field(0, t), SpecialForm("WHEN", BIGINT, [field(1, t), constant(1, BIGINT)]), SpecialForm("WHEN", BIGINT, [field(2, t), constant(2, BIGINT)])]
        if with_else:
            args.append(constant(0, BIGINT))
        plan = plan_text([t, t, t], None, [SpecialForm("SWITCH", BIGINT, args)])
        exp = [(simple_case(a, b, c, 0 if with_else else None),) for a, b, c in zip(v, f, s2)]
        assert {r[0] for r in exp} >= {1, 2}
        cases.append(("simple_case_grid_" + name, plan, [Page(block_of(t, v), block_of(t, f), block_of(t, s2))], exp, "ordered"))

    # TestConditions.testSimpleCase literals on a BOOLEAN value (one row each, the value comes from the page)
    T, F, N = constant(True, BOOLEAN), constant(False, BOOLEAN), constant(None, BOOLEAN)
    V = field(0, BOOLEAN)

    def when(o, r):
        return SpecialForm("WHEN", BIGINT, [o, r])

    c33, c1, cnull = constant(33, BIGINT), constant(1, BIGINT), constant(None, BIGINT)
    proj = [SpecialForm("SWITCH", BIGINT, [V, when(T, c33)]),                                  # :318  -> 33
            SpecialForm("SWITCH", BIGINT, [V, when(F, c1), c33]),                              # :332  -> 33
            SpecialForm("SWITCH", BIGINT, [V, when(F, c1), when(F, c1), when(T, c33), c1]),    # :346  -> 33
            SpecialForm("SWITCH", BIGINT, [V, when(F, c1)]),                                   # :355  -> NULL
            SpecialForm("SWITCH", BIGINT, [V, when(T, cnull), c1]),                            # :361  -> NULL (a NULL result is a result)
            SpecialForm("SWITCH", BIGINT, [V, when(N, c1), when(T, c33)]),                     # :369  -> 33 (NULL operand never matches)
            SpecialForm("SWITCH", BIGINT, [constant(None, BOOLEAN), when(T, c1), c33]),        # :383  -> 33 (NULL value takes ELSE)
            SpecialForm("SWITCH", BIGINT, [V, when(V, c33), c1])]                              # value = operand, NULL = NULL does not match
    plan = plan_text([BOOLEAN], None, proj)
    exp = [(33, 33, 33, None, None, 33, 33, 33), (None, 1, 1, 1, 1, None, 33, 33), (None, 33, 1, None, 1, None, 33, 1)]
    cases.append(("simple_case_literals", plan, [Page(block_of(BOOLEAN, [True, False, None]))], exp, "ordered"))

    # CASE and NULLIF feeding a filter and an aggregation: sum(nullif(x, 0)), count(case k when 1 then x when 2 then -x end) by k
    n = 4000
    k = [None if i % 17 == 0 else i % 4 for i in range(n)]
    x = [None if i % 19 == 0 else (i * 7) % 11 - 5 for i in range(n)]
    K, X = field(0, BIGINT), field(1, BIGINT)
    sw = SpecialForm("SWITCH", BIGINT, [K, SpecialForm("WHEN", BIGINT, [constant(1, BIGINT), X]), SpecialForm("WHEN", BIGINT, [constant(2, BIGINT), call("negate", BIGINT, X)])])
    f = compare("ne", SpecialForm("COALESCE", BIGINT, [SpecialForm("NULL_IF", BIGINT, [X, constant(3, BIGINT)]), constant(0, BIGINT)]), constant(-4, BIGINT))
    plan = plan_text([BIGINT, BIGINT], f, [K, SpecialForm("NULL_IF", BIGINT, [X, constant(0, BIGINT)]), sw], step="single", groupByChannels=[0],
                     aggregations=[Aggregation("sum", BIGINT, 1, BIGINT), Aggregation("count", BIGINT, 2, BIGINT), Aggregation("sum", BIGINT, 2, BIGINT)], expectedGroups=16)
    groups = {}
    for a, b in zip(k, x):
        if (0 if (b is None or b == 3) else b) == -4:
            continue
        g = groups.setdefault(a, [[], []])
        if b is not None and b != 0:
            g[0].append(b)
        c = None if a is None else (b if a == 1 else (None if b is None else -b) if a == 2 else None)
        if c is not None:
            g[1].append(c)
    exp = [(a, sum(g[0]) if g[0] else None, len(g[1]), sum(g[1]) if g[1] else None) for a, g in groups.items()]
    cases.append(("case_nullif_into_aggregation", plan, [Page(block_of(BIGINT, k), block_of(BIGINT, x))], exp, 1))
    return cases


def operator_test_cases():
    """TestAggregationOperator.java:91-118 (mask with dirty nulls), :120-156 (global aggregation over sequence pages, the functions of
    the supported subset), TestFilterAndProjectOperator.java:80-124 (filter bigint <= 9, project the varchar and bigint + 5)."""
    from presto_b200.spi import ByteArrayBlock, VARCHAR
    cases = []
    # count(channel 0) masked by a BOOLEAN channel whose NULL positions hold garbage and whose TRUE is any non-zero byte
    mask = ByteArrayBlock(BOOLEAN, np.array([0, 27, 0, 75], dtype=np.int8), np.array([1, 1, 0, 0], dtype=np.uint8))
    plan = plan_text([BIGINT, BOOLEAN], step="single", aggregations=[Aggregation("count", BIGINT, 0, BIGINT, maskChannel=1)])
    cases.append(("optest_mask_with_dirty_nulls", plan, [Page(block_of(BIGINT, [1, 2, 3, 4]), mask)], [(1,)], None))
    # addSequencePage(100, 0, 0, 300, 500, 400, 500, 500): channel i holds start_i .. start_i + 99 (varchar: the decimal string)
    v10 = createVarcharType(10)
    seq = lambda start: list(range(start, start + 100))
    page = Page(block_of(v10, [str(x) for x in seq(0)]), block_of(BIGINT, seq(0)), block_of(BIGINT, seq(500)), block_of(DOUBLE, [float(x) for x in seq(500)]))
    plan = plan_text([v10, BIGINT, BIGINT, DOUBLE], step="single",
                     aggregations=[Aggregation("count", BIGINT, 0, v10), Aggregation("sum", BIGINT, 1, BIGINT), Aggregation("avg", DOUBLE, 1, BIGINT),
                                   Aggregation("count", BIGINT, 0, v10), Aggregation("sum", BIGINT, 2, BIGINT), Aggregation("sum", DOUBLE, 3, DOUBLE)])
    cases.append(("optest_global_aggregation", plan, [page], [(100, 4950, 49.5, 100, 54950, 54950.0)], None))
    # FilterAndProjectOperator
    page = Page(block_of(v10, [str(x) for x in seq(0)]), block_of(BIGINT, seq(0)))
    plan = plan_text([v10, BIGINT], compare("le", field(1, BIGINT), constant(9, BIGINT)), [field(0, v10), call("add", BIGINT, field(1, BIGINT), constant(5, BIGINT))])
    cases.append(("optest_filter_and_project", plan, [page], [(str(i), i + 5) for i in range(10)], "ordered"))
    return cases


def dictionary_rle_cases():
    """TestDictionaryAwarePageFilter (:49-116) / TestDictionaryAwarePageProjection (:80-143): dictionary, RLE and plain
    inputs holding the same values must give the same result."""
    cases = []
    V3 = createVarcharType(3)
    n = 5000
    dict_vals = ["a", "bb", None, "ccc", ""]
    ids = [(i * 7) % 5 for i in range(n)]
    price = [float(i % 100) for i in range(n)]
    dblock = VariableWidthBlock.from_values(V3, dict_vals)
    plain = VariableWidthBlock.from_values(V3, [dict_vals[i] for i in ids])
    qty_dict = LongArrayBlock(DOUBLE, np.arange(1, 51, dtype=np.float64))
    qty_ids = [(i * 13) % 50 for i in range(n)]
    qty_plain = LongArrayBlock(DOUBLE, np.array([1.0 + q for q in qty_ids]))
    plan = plan_text([V3, DOUBLE, DOUBLE, BIGINT], compare("lt", field(2, DOUBLE), constant(40.0, DOUBLE)),
                     [field(0, V3), call("multiply", DOUBLE, field(1, DOUBLE), field(2, DOUBLE)), field(3, BIGINT)],
                     step="single", groupByChannels=[0, 2],
                     aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE), Aggregation("count", BIGINT), Aggregation("min", DOUBLE, 1, DOUBLE)])
    groups = {}
    for i in range(n):
        q = 1.0 + qty_ids[i]
        if q < 40.0:
            g = groups.setdefault(dict_vals[ids[i]], [])
            g.append(price[i] * q)
    exp = [(k, 42, math.fsum(v), len(v), min(v)) for k, v in groups.items()]
    rle = RunLengthEncodedBlock(block_of(BIGINT, [42]), n)
    enc = Page(DictionaryBlock(dblock, ids), block_of(DOUBLE, price), DictionaryBlock(qty_dict, qty_ids), rle)
    flat = Page(plain, block_of(DOUBLE, price), qty_plain, block_of(BIGINT, [42] * n))
    cases.append(("dictionary_rle_inputs", plan, [enc], exp, 2))
    cases.append(("dictionary_rle_inputs_as_plain", plan, [flat], exp, 2))
    cases.append(("dictionary_then_plain_pages", plan, [enc, flat], [(k, c, 2 * s, 2 * cnt, m) for k, c, s, cnt, m in exp], 2))
    # RLE NULL column: every aggregate sees only NULLs (AbstractTestAggregationFunction.testAllPositionsNull uses RLE nulls)
    plan = plan_text([DOUBLE], step="single", aggregations=[Aggregation("sum", DOUBLE, 0, DOUBLE), Aggregation("count", BIGINT, 0, DOUBLE), Aggregation("count", BIGINT)])
    cases.append(("rle_null_column", plan, [Page(RunLengthEncodedBlock(block_of(DOUBLE, [None]), 10))], [(None, 0, 10)], None))
    return cases


def all_cases():
    return global_aggregate_cases() + hash_aggregation_cases() + filter_project_cases() + conditional_cases() + operator_test_cases() + dictionary_rle_cases()
