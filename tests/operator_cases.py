"""Scalar operator known-answer vectors restated from the reference's own operator tests (paths under
/root/reference/core/trino-main/src/test/java/io/trino/type/): TestBigintOperators.java:60-162,236-272, TestIntegerOperators.java:
60-115, TestDoubleOperators.java:66-190,237-260, TestDecimalOperators.java:32-69,141-177 — the semantics of
main/type/{Bigint,Integer,Double,Decimal}Operators.java that the plan's `call` nodes must reproduce (§8 A6).

value_cases():  (name, plan, pages, expected rows, "ordered") like tests/kat_cases.py
error_cases():  (name, plan, pages, StandardErrorCode name)
The operands travel as columns (one row per assertion of the reference test), so every operator is evaluated by the kernel, never
folded at plan time."""
import math

from presto_b200.operators import plan_text
from presto_b200.relational import SpecialForm, call, compare, constant, field
from presto_b200.spi import BIGINT, BOOLEAN, DOUBLE, INTEGER, Page, block_of, createDecimalType

NAN, INF = float("nan"), float("inf")
I64_MAX, I64_MIN, I32_MAX, I32_MIN = 2 ** 63 - 1, -2 ** 63, 2 ** 31 - 1, -2 ** 31


def _jdiv(a, b):      # Java integer division truncates toward zero
    q = abs(a) // abs(b)
    return q if (a < 0) == (b < 0) else -q


def _jmod(a, b):      # remainder has the dividend's sign
    return a - _jdiv(a, b) * b


def _cmp_projections(t, A, B):
    return [compare(op, A, B) for op in ("eq", "ne", "lt", "le", "gt", "ge")] + [SpecialForm("BETWEEN", BOOLEAN, [A, B, A])]


def value_cases():
    cases = []
    # ---- BIGINT (TestBigintOperators) and INTEGER (TestIntegerOperators): same operand pairs, the INTEGER ones scaled to 32 bits
    for name, t, big, other in (("bigint", BIGINT, 100000000037, 100000000017), ("integer", INTEGER, 37, 17)):
        small, mulb = (37, 10000017) if t is BIGINT else (37, 17)
        pairs = [(big, small), (small, other), (other, small), (other, other), (big, big), (big, other), (other, big), (-big, small), (small, -other), (-other, -small)]
        a, b = [p[0] for p in pairs], [p[1] for p in pairs]
        A, B = field(0, t), field(1, t)
        m = constant(mulb, t)
        proj = [call("add", t, A, B), call("subtract", t, A, B), call("multiply", t, A, m), call("divide", t, A, B), call("modulus", t, A, B), call("negate", t, A)]
        proj += _cmp_projections(t, A, B)
        exp = [(x + y, x - y, x * mulb, _jdiv(x, y), _jmod(x, y), -x, x == y, x != y, x < y, x <= y, x > y, x >= y, y <= x <= x) for x, y in pairs]
        if t is INTEGER:
            assert all(I32_MIN <= v <= I32_MAX for r in exp for v in r[:6])
        cases.append((name + "_operators", plan_text([t, t], None, proj), [Page(block_of(t, a), block_of(t, b))], exp, "ordered"))
    # extremes that must NOT overflow
    t = BIGINT
    a, b = [I64_MAX, I64_MIN, I64_MIN + 1, I64_MAX, I64_MIN], [0, 0, 1, -1, 1]
    A, B = field(0, t), field(1, t)
    proj = [call("add", t, A, B), call("multiply", t, A, B), call("divide", t, A, constant(2, t)), call("modulus", t, A, constant(-1, t))]
    exp = [(x + y, x * y, _jdiv(x, 2), 0) for x, y in zip(a, b)]
    exp[4] = (I64_MIN + 1, I64_MIN, _jdiv(I64_MIN, 2), 0)
    assert all(I64_MIN <= v <= I64_MAX for r in exp for v in r)
    cases.append(("bigint_extremes", plan_text([t, t], None, proj), [Page(block_of(t, a), block_of(t, b))], exp, "ordered"))

    # ---- DOUBLE (TestDoubleOperators): NaN propagates through arithmetic, every comparison with NaN is false except <>
    vals = [(37.7, 37.7), (37.7, 17.1), (17.1, 37.7), (17.1, 17.1), (NAN, 37.7), (37.7, NAN), (NAN, NAN), (-37.7, 17.1), (0.0, -0.0), (INF, 1.0), (1.0, 0.0), (-1.0, 0.0)]
    a, b = [p[0] for p in vals], [p[1] for p in vals]
    A, B = field(0, DOUBLE), field(1, DOUBLE)
    proj = [call("add", DOUBLE, A, B), call("subtract", DOUBLE, A, B), call("multiply", DOUBLE, A, B), call("divide", DOUBLE, A, B), call("modulus", DOUBLE, A, B),
            call("negate", DOUBLE, A)] + [compare(op, A, B) for op in ("eq", "ne", "lt", "le", "gt", "ge")]

    def ddiv(x, y):
        if y == 0.0 and not math.isnan(x):
            return NAN if x == 0.0 else math.copysign(INF, x) * math.copysign(1.0, y)
        return x / y

    def dmod(x, y):  # Java %, = C fmod
        if math.isnan(x) or math.isnan(y) or math.isinf(x) or y == 0.0:
            return NAN
        return math.fmod(x, y)
    exp = [(x + y, x - y, x * y, ddiv(x, y), dmod(x, y), -x, x == y, x != y, x < y, x <= y, x > y, x >= y) for x, y in vals]
    cases.append(("double_operators", plan_text([DOUBLE, DOUBLE], None, proj), [Page(block_of(DOUBLE, a), block_of(DOUBLE, b))], exp, "ordered"))
    # casts (TestDoubleOperators.java:237-253: round half away from zero; TestBigintOperators.java:196-199)
    d = [37.7, -37.7, 17.1, -17.1, 9.2e18, -9.2e18, 2.21e9, -2.21e9, 17.5, -17.5, math.nextafter(2.0 ** 63, 0.0), -2.0 ** 63, math.nextafter(-2.0 ** 63, 0.0), 0.49999999999999994, -0.5]
    want = [38, -38, 17, -17, 9200000000000000000, -9200000000000000000, 2210000000, -2210000000, 18, -18, int(math.nextafter(2.0 ** 63, 0.0)), -2 ** 63,
            int(math.nextafter(-2.0 ** 63, 0.0)), 0, -1]
    i = [37, 100000000017, -100000000017, 0, 2 ** 53 + 1, I64_MAX, I64_MIN] + [0] * 8
    proj = [call("cast", BIGINT, field(0, DOUBLE)), call("cast", DOUBLE, field(1, BIGINT))]
    exp = [(w, float(x)) for w, x in zip(want, i)]
    cases.append(("double_bigint_casts", plan_text([DOUBLE, BIGINT], None, proj), [Page(block_of(DOUBLE, d), block_of(BIGINT, i))], exp, "ordered"))

    # ---- DECIMAL (TestDecimalOperators.java:32-69 add, :141-167 multiply): literal -> type by its digits, result type by the rules of
    # DecimalOperators.java:84-85,261-262; one plan per assertion because every assertion has its own types
    def lit(s):
        neg = s.startswith("-")
        body = s.lstrip("+-")
        whole, _, frac = body.partition(".")
        digits = (whole + frac).lstrip("0") or "0"
        precision = max(len(whole + frac) if whole else len(frac), 1)
        v = int(digits)
        return createDecimalType(precision, len(frac)), -v if neg else v

    def result_type(op, ta, tb):
        if op == "add":
            s = max(ta.scale, tb.scale)
            return createDecimalType(min(38, max(ta.precision - ta.scale, tb.precision - tb.scale) + s + 1), s)
        return createDecimalType(min(38, ta.precision + tb.precision), ta.scale + tb.scale)

    vectors = [("add", "137.7", "17.1", "0154.8"), ("add", "-1", "-2", "-03"), ("add", "1", "2", "03"),
               ("add", ".1234567890123456", ".1234567890123456", "0.2469135780246912"), ("add", "-.1234567890123456", "-.1234567890123456", "-0.2469135780246912"),
               ("add", "1234567890123456", "1234567890123456", "02469135780246912"), ("add", "123456789012345678", "123456789012345678", "0246913578024691356"),
               ("add", ".123456789012345678", ".123456789012345678", "0.246913578024691356"), ("add", "1234567890123456789", "1234567890123456789", "02469135780246913578"),
               ("add", "12345678901234567890123456789012345678", "12345678901234567890123456789012345678", "24691357802469135780246913578024691356"),
               ("add", "-99999999999999999999999999999999999999", "99999999999999999999999999999999999999", "00000000000000000000000000000000000000"),
               ("add", "12345678901234567890", "-12345678901234567890", "000000000000000000000"), ("add", "-12345678901234567890", "-12345678901234567890", "-024691357802469135780"),
               ("add", "12345678901234567890", "-12345678901234567891", "-000000000000000000001"), ("add", "-12345678901234567890", "12345678901234567891", "000000000000000000001"),
               ("add", "999999999999999999", "999999999999999999", "1999999999999999998"), ("add", "999999999999999999", ".999999999999999999", "0999999999999999999.999999999999999999"),
               ("add", "123456789012345678901234567890", ".12345678", "123456789012345678901234567890.12345678"),
               ("add", ".12345678", "123456789012345678901234567890", "123456789012345678901234567890.12345678"),
               ("multiply", "12", "3", "036"), ("multiply", "12", "-3", "-036"), ("multiply", "-12", "3", "-036"), ("multiply", "1234567890123456", "3", "03703703670370368"),
               ("multiply", ".1234567890123456", "3", "0.3703703670370368"), ("multiply", ".1234567890123456", ".3", ".03703703670370368"),
               ("multiply", "12345678901234567", "123456789012345670", "01524157875323883455265967556774890"),
               ("multiply", "-12345678901234567", "123456789012345670", "-01524157875323883455265967556774890"),
               ("multiply", "-12345678901234567", "-123456789012345670", "01524157875323883455265967556774890"),
               ("multiply", ".12345678901234567", ".123456789012345670", ".01524157875323883455265967556774890"),
               ("multiply", "12345678901234567890123456789012345678", "3", "37037036703703703670370370367037037034"),
               ("multiply", "1234567890123456789.0123456789012345678", "3", "3703703670370370367.0370370367037037034"),
               ("multiply", "3", ".12345678901234567890123456789012345678", ".37037036703703703670370370367037037034"),
               ("multiply", ".1234567890123456789", ".1234567890123456789", ".01524157875323883675019051998750190521")]
    for k, (op, sa, sb, sr) in enumerate(vectors):
        (ta, va), (tb, vb), (tr, vr) = lit(sa), lit(sb), lit(sr)
        rt = result_type(op, ta, tb)
        assert (rt.precision, rt.scale) == (tr.precision, tr.scale), (op, sa, sb, sr, rt, tr)
        plan = plan_text([ta, tb], None, [call(op, rt, field(0, ta), field(1, tb)), call(op, rt, field(0, ta), constant(vb, tb))])
        cases.append(("decimal_%s_%02d" % (op, k), plan, [Page(block_of(ta, [va, None]), block_of(tb, [vb, vb]))], [(vr, vr), (None, None)], "ordered"))
    return cases


def error_cases():
    cases = []

    def one(name, t, proj, a, b, code):
        cases.append((name, plan_text([t, t], None, [proj]), [Page(block_of(t, [1, a]), block_of(t, [1, b]))], code))
    for name, t, hi, lo in (("bigint", BIGINT, I64_MAX, I64_MIN), ("integer", INTEGER, I32_MAX, I32_MIN)):
        A, B = field(0, t), field(1, t)
        ovf = "NUMERIC_VALUE_OUT_OF_RANGE"
        one(name + "_add_overflow", t, call("add", t, A, B), hi, 1, ovf)                 # TestBigintOperators.java:236-239
        one(name + "_subtract_underflow", t, call("subtract", t, A, B), lo, 1, ovf)      # :242-245
        one(name + "_multiply_overflow", t, call("multiply", t, A, B), hi, 2, ovf)       # :248-250
        one(name + "_multiply_min_by_minus_one", t, call("multiply", t, A, B), lo, -1, ovf)  # :251
        one(name + "_divide_overflow", t, call("divide", t, A, B), lo, -1, ovf)          # :255-258
        one(name + "_negate_overflow", t, call("negate", t, A), lo, 1, ovf)              # :270-273
        one(name + "_divide_by_zero", t, call("divide", t, A, B), 17, 0, "DIVISION_BY_ZERO")   # TestIntegerOperators.java:95
        one(name + "_modulus_by_zero", t, call("modulus", t, A, B), 17, 0, "DIVISION_BY_ZERO")  # :105
    for k, v in enumerate([2.0 ** 63, math.nextafter(2.0 ** 63, INF), math.nextafter(-2.0 ** 63, -INF), 9.3e18, -9.3e18, INF, -INF, NAN]):   # TestDoubleOperators.java:249-260
        cases.append(("double_to_bigint_invalid_%d" % k, plan_text([DOUBLE], None, [call("cast", BIGINT, field(0, DOUBLE))]), [Page(block_of(DOUBLE, [1.0, v]))], "INVALID_CAST_ARGUMENT"))
    d38, d1, d38_38 = createDecimalType(38, 0), createDecimalType(1, 0), createDecimalType(38, 38)
    nines = 10 ** 38 - 1
    big = 12345678901234567890123456789012345678
    for name, ta, tb, op, rt, a, b in (("decimal_add_overflow", d38, d1, "add", d38, nines, 1),                       # TestDecimalOperators.java:68
                                       ("decimal_multiply_overflow", d38, d1, "multiply", d38, big, 9),               # :174
                                       ("decimal_multiply_overflow_scaled", d38_38, d1, "multiply", d38_38, big, 9),  # :175
                                       ("decimal_multiply_overflow_negative", d38, d1, "multiply", d38, big, -9)):    # :176
        cases.append((name, plan_text([ta, tb], None, [call(op, rt, field(0, ta), field(1, tb))]), [Page(block_of(ta, [1, a]), block_of(tb, [1, b]))], "NUMERIC_VALUE_OUT_OF_RANGE"))
    return cases
