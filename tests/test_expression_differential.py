"""Differential test of the product's generated expression code (executed on the CPU from the GPU source, tests/emu/expr_emu.py)
against the CPU oracle, beyond the reference's literal vectors: random operands in bulk, and every pair of boundary values one row at
a time, where the oracle either returns a value or raises and the generated code must return the same value or raise the matching
device error bit."""
import itertools
import math
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))
import expr_emu  # noqa: E402
from presto_b200.operators import GpuPipelineOperator, TrinoException, plan_text  # noqa: E402
from presto_b200.relational import SpecialForm, call, compare, constant, field  # noqa: E402
from presto_b200.spi import BIGINT, BOOLEAN, DOUBLE, INTEGER, Page, block_of, createDecimalType  # noqa: E402
from test_expression_emu import ERR_BITS, execute  # noqa: E402
from util import run_oracle, value_equal  # noqa: E402

I64_MAX, I64_MIN, I32_MAX, I32_MIN = 2 ** 63 - 1, -2 ** 63, 2 ** 31 - 1, -2 ** 31


def binary_plans(t, result=None):
    r = result or (lambda op: t)
    A, B = field(0, t), field(1, t)
    plans = {op: plan_text([t, t], None, [call(op, r(op), A, B)]) for op in ("add", "subtract", "multiply", "divide", "modulus")}
    plans["negate"] = plan_text([t, t], None, [call("negate", t, A)])
    plans["compare"] = plan_text([t, t], None, [compare(op, A, B) for op in ("eq", "ne", "lt", "le", "gt", "ge")] + [SpecialForm("BETWEEN", BOOLEAN, [A, B, A]),
                                                 SpecialForm("NULL_IF", t, [A, B]), SpecialForm("COALESCE", t, [A, B])])
    return plans


def same_rows(got, want):
    assert len(got) == len(want)
    for i, (g, w) in enumerate(zip(got, want)):
        assert len(g) == len(w) and all(value_equal(x, y, 0.0) and (not isinstance(x, float) or x != 0.0 or math.copysign(1, x) == math.copysign(1, y)) for x, y in zip(g, w)), (i, g, w)


def check_row_by_row(plan, t, pairs):
    """each pair alone in a page: value or fault, the same on both sides"""
    outcomes = set()
    for a, b in pairs:
        page = [Page(block_of(t, [a]), block_of(t, [b]))]
        try:
            want, code = run_oracle(plan, page), None
        except TrinoException as e:
            want, code = None, e.code
        got, err = execute(plan, page)
        if code is None:
            assert err == 0, (a, b, err)
            same_rows(got, want)
        else:
            assert err & ERR_BITS[code], (a, b, code, err)
        outcomes.add(code)
    return outcomes


@pytest.mark.parametrize("name,t,lo,hi", [("bigint", BIGINT, I64_MIN, I64_MAX), ("integer", INTEGER, I32_MIN, I32_MAX)])
def test_integer_operators_at_the_boundaries(name, t, lo, hi):
    edge = [lo, lo + 1, -2, -1, 0, 1, 2, hi - 1, hi, None]
    pairs = list(itertools.product(edge, edge))
    seen = set()
    for op, plan in binary_plans(t).items():
        seen |= check_row_by_row(plan, t, pairs)
    assert {None, "NUMERIC_VALUE_OUT_OF_RANGE", "DIVISION_BY_ZERO"} <= seen


def test_integer_operators_random_bulk():
    rng = np.random.default_rng(17)
    n = 4000
    for t, span in ((BIGINT, 2 ** 31), (INTEGER, 2 ** 15)):   # products stay in range
        a = [None if rng.random() < 0.05 else int(v) for v in rng.integers(-span, span, n)]
        b = [None if rng.random() < 0.05 else int(v) or 1 for v in rng.integers(-span, span, n)]
        for op, plan in binary_plans(t).items():
            page = [Page(block_of(t, a), block_of(t, b))]
            got, err = execute(plan, page)
            assert err == 0
            same_rows(got, run_oracle(plan, page))


def test_double_operators_specials_and_random():
    specials = [0.0, -0.0, 1.0, -1.0, 0.1, 5e-324, -5e-324, 2.2250738585072014e-308, 1.7976931348623157e308, -1.7976931348623157e308, float("inf"), float("-inf"),
                float("nan"), 2.0 ** 53, 2.0 ** 63, -2.0 ** 63, 1e308, 37.7, -17.1, None]
    pairs = list(itertools.product(specials, specials))
    a, b = [p[0] for p in pairs], [p[1] for p in pairs]
    rng = np.random.default_rng(3)
    a += list(rng.normal(size=3000) * 10.0 ** rng.integers(-300, 300, 3000))
    b += list(rng.normal(size=3000) * 10.0 ** rng.integers(-300, 300, 3000))
    for op, plan in binary_plans(DOUBLE).items():
        page = [Page(block_of(DOUBLE, a), block_of(DOUBLE, b))]
        got, err = execute(plan, page)
        assert err == 0
        same_rows(got, run_oracle(plan, page))   # bit-exact per row: Java has no FMA contraction, the kernels are built with --fmad=false
    cast = plan_text([DOUBLE, DOUBLE], None, [call("cast", BIGINT, field(0, DOUBLE))])
    seen = check_row_by_row(cast, DOUBLE, [(x, 0.0) for x in specials + [0.5, -0.5, 1.5, 2.5, -2.5, 0.49999999999999994, 9.223372036854775e18, -9.223372036854777e18, 4503599627370497.5]])
    assert seen == {None, "INVALID_CAST_ARGUMENT"}


def test_decimal_operators_boundaries_and_random():
    """short x short -> short and -> long, long x long: add / subtract / multiply with the reference's result types; overflow past 38 digits"""
    d12, d18, d30, d38 = createDecimalType(12, 2), createDecimalType(18, 0), createDecimalType(30, 4), createDecimalType(38, 0)

    def result(ta, tb):
        def r(op):
            if op in ("add", "subtract"):
                s = max(ta.scale, tb.scale)
                return createDecimalType(min(38, max(ta.precision - ta.scale, tb.precision - tb.scale) + s + 1), s)
            return createDecimalType(min(38, ta.precision + tb.precision), ta.scale + tb.scale)
        return r
    import random
    rng, pyrng = np.random.default_rng(8), random.Random(8)
    for t in (d12, d18, d30, d38):
        top = 10 ** t.precision - 1
        edge = [-top, -top + 1, -10 ** (t.precision // 2), -1, 0, 1, 10 ** (t.precision // 2), top - 1, top, None]
        bulk = [int(rng.integers(-10 ** 9, 10 ** 9)) * pyrng.randrange(1, 10 ** (t.precision - 9)) for _ in range(300)]
        A, B = field(0, t), field(1, t)
        r = result(t, t)
        plans = {op: plan_text([t, t], None, [call(op, r(op), A, B)]) for op in ("add", "subtract", "multiply")}
        plans["negate"] = plan_text([t, t], None, [call("negate", t, A)])
        plans["compare"] = plan_text([t, t], None, [compare(op, A, B) for op in ("eq", "ne", "lt", "le", "gt", "ge")])
        seen = set()
        for op, plan in plans.items():
            seen |= check_row_by_row(plan, t, list(itertools.product(edge, edge)))
            page = [Page(block_of(t, bulk), block_of(t, bulk[::-1]))]
            try:
                want = run_oracle(plan, page)
            except TrinoException:
                continue   # an overflowing product somewhere in the bulk: covered row by row above
            got, err = execute(plan, page)
            assert err == 0
            same_rows(got, want)
        if t.precision >= 30:
            assert "NUMERIC_VALUE_OUT_OF_RANGE" in seen


def test_a_null_argument_stops_the_evaluation_of_the_later_ones():
    """BytecodeUtils.generateFullInvocation: x + (y / z) with x NULL never evaluates y / z (no DIVISION_BY_ZERO); with x present and
    z = 0 it raises; a fault under a NULL in the same call (NULL - MIN) is not raised either"""
    A, B, C = field(0, BIGINT), field(1, BIGINT), field(2, BIGINT)
    plan = plan_text([BIGINT, BIGINT, BIGINT], None, [call("add", BIGINT, A, call("divide", BIGINT, B, C)), call("subtract", BIGINT, A, B)])

    def both(a, b, c):
        page = [Page(block_of(BIGINT, [a]), block_of(BIGINT, [b]), block_of(BIGINT, [c]))]
        try:
            want = run_oracle(plan, page)
        except TrinoException as e:
            want = e.code
        got, err = execute(plan, page)
        return want, got, err
    want, got, err = both(None, 10, 0)
    assert want == [(None, None)] and got == want and err == 0
    want, got, err = both(None, I64_MIN, 3)            # NULL - MIN: no overflow
    assert want == [(None, None)] and got == want and err == 0
    want, got, err = both(1, 10, 0)
    assert want == "DIVISION_BY_ZERO" and err & ERR_BITS["DIVISION_BY_ZERO"]
    want, got, err = both(7, 10, 3)
    assert want == [(10, -3)] and got == want and err == 0
    # dirty values under NULL positions must not matter either
    from presto_b200.spi import LongArrayBlock
    dirty = LongArrayBlock(BIGINT, np.array([I64_MAX, 5], dtype=np.int64), np.array([1, 0], dtype=np.uint8))
    page = [Page(dirty, block_of(BIGINT, [1, 1]), block_of(BIGINT, [1, 1]))]
    got, err = execute(plan, page)
    assert err == 0 and got == run_oracle(plan, page) == [(None, None), (6, 4)]


def test_casts_at_the_boundaries():
    """every cast the plan language accepts between BIGINT / INTEGER / DOUBLE / BOOLEAN / DECIMAL(p,s), over boundary values one row at
    a time: same value (rounding included) or matching fault on both sides"""
    d5_2, d12_2, d18_0, d20_4, d38_10 = (createDecimalType(*x) for x in ((5, 2), (12, 2), (18, 0), (20, 4), (38, 10)))
    types = [BIGINT, INTEGER, DOUBLE, BOOLEAN, d5_2, d12_2, d18_0, d20_4, d38_10]

    def samples(t):
        if t is BIGINT:
            return [I64_MIN, I64_MIN + 1, -10 ** 18, I32_MIN - 1, I32_MIN, -100000, -1, 0, 1, 99999, I32_MAX, I32_MAX + 1, 10 ** 18, I64_MAX, None]
        if t is INTEGER:
            return [I32_MIN, -100000, -1, 0, 1, 999, 99999, I32_MAX, None]
        if t is DOUBLE:
            return [0.0, -0.0, 0.5, -0.5, 1.5, 2.5, 0.005, 0.015, 999.995, 999.994999, -999.995, 1e10, -1e10, 1e19, 1e38, 1e39, float("nan"), float("inf"), 123456.789, None]
        if t is BOOLEAN:
            return [True, False, None]
        top = 10 ** t.precision - 1
        half = 5 * 10 ** max(t.scale - 1, 0)
        return [-top, -top + 1, -half, -half + 1, -1, 0, 1, half - 1, half, half + 1, 10 ** t.scale, 15 * 10 ** max(t.scale - 1, 0), top - 1, top, None]
    supported, outcomes = 0, set()
    for f in types:
        for to in types:
            if f == to:
                continue
            plan = plan_text([f, f], None, [call("cast", to, field(0, f))])
            try:
                GpuPipelineOperator(plan, device=-1).close()
            except TrinoException as e:
                assert e.code == "NOT_SUPPORTED", (f, to, e)
                continue
            supported += 1
            outcomes |= check_row_by_row(plan, f, [(v, v) for v in samples(f)])
    assert supported >= 20 and {None, "NUMERIC_VALUE_OUT_OF_RANGE", "INVALID_CAST_ARGUMENT"} <= outcomes


def test_short_circuit_forms_do_not_raise_in_untaken_branches():
    """AND / OR (AndCodeGenerator / OrCodeGenerator stop at the first FALSE / TRUE), IF and COALESCE evaluate only what they need: a
    division by zero in an operand that is not reached is not raised; where it is reached, it is — identically in the oracle and in
    the generated code"""
    X, Y = field(0, BIGINT), field(1, BIGINT)
    one, zero = constant(1, BIGINT), constant(0, BIGINT)
    faulty = compare("eq", call("divide", BIGINT, one, Y), one)             # 1 / y = 1
    from presto_b200.relational import and_, or_
    plans = {"and": plan_text([BIGINT, BIGINT], None, [and_(compare("lt", X, zero), faulty)]),
             "or": plan_text([BIGINT, BIGINT], None, [or_(compare("ge", X, zero), faulty)]),
             "if": plan_text([BIGINT, BIGINT], None, [SpecialForm("IF", BIGINT, [compare("lt", X, zero), call("divide", BIGINT, one, Y), X])]),
             "coalesce": plan_text([BIGINT, BIGINT], None, [SpecialForm("COALESCE", BIGINT, [X, call("divide", BIGINT, one, Y)])]),
             "filter": plan_text([BIGINT, BIGINT], and_(compare("lt", X, zero), faulty), [X])}
    pairs = [(5, 0), (-5, 0), (5, 1), (-5, 1), (None, 0), (None, 1), (0, None), (-1, None)]
    for name, plan in plans.items():
        seen = check_row_by_row(plan, BIGINT, pairs)
        assert seen == {None, "DIVISION_BY_ZERO"}, (name, seen)
    # and the unreached operand really is unreached: x >= 0 with y = 0 gives FALSE / TRUE / x / x without a fault
    for name, want in (("and", [(False,)]), ("or", [(True,)]), ("if", [(5,)]), ("coalesce", [(5,)]), ("filter", [])):
        got, err = execute(plans[name], [Page(block_of(BIGINT, [5]), block_of(BIGINT, [0]))])
        assert err == 0 and got == want, name


def test_between_evaluates_its_bounds_like_the_reference():
    """BetweenCodeGenerator: value >= min AND value <= max. A faulting upper bound is not reached where value < min, neither bound where
    the value is NULL; where they are reached the fault is raised — oracle and generated code alike"""
    V, Y = field(0, BIGINT), field(1, BIGINT)
    ten = constant(10, BIGINT)
    hundred_over_y = call("divide", BIGINT, constant(100, BIGINT), Y)
    upper = plan_text([BIGINT, BIGINT], None, [SpecialForm("BETWEEN", BOOLEAN, [V, ten, hundred_over_y])])
    lower = plan_text([BIGINT, BIGINT], None, [SpecialForm("BETWEEN", BOOLEAN, [V, hundred_over_y, ten])])
    pairs = [(5, 0), (15, 0), (15, 2), (60, 2), (None, 0), (None, 2), (15, None), (5, None), (10, 10), (10, 0)]
    for plan in (upper, lower):
        assert check_row_by_row(plan, BIGINT, pairs) == {None, "DIVISION_BY_ZERO"}
    for plan, row, want in ((upper, (5, 0), [(False,)]), (upper, (None, 0), [(None,)]), (lower, (None, 0), [(None,)]), (upper, (15, None), [(None,)]), (upper, (5, None), [(False,)])):
        page = [Page(block_of(BIGINT, [row[0]]), block_of(BIGINT, [row[1]]))]
        got, err = execute(plan, page)
        assert err == 0 and got == want == run_oracle(plan, page), (row, got)
    for plan, row in ((upper, (15, 0)), (lower, (15, 0))):
        page = [Page(block_of(BIGINT, [row[0]]), block_of(BIGINT, [row[1]]))]
        with pytest.raises(TrinoException):
            run_oracle(plan, page)
        assert execute(plan, page)[1] & ERR_BITS["DIVISION_BY_ZERO"]
