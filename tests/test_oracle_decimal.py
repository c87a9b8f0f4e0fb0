"""Pins oracle/decimal128.h against the literal vectors of the reference's
core/trino-spi/src/test/java/io/trino/spi/type/TestUnscaledDecimal128Arithmetic.java (:101-205, :294-330) and the
state vectors of core/trino-main/src/test/java/io/trino/operator/aggregation/TestDecimalSumAggregation.java:46-124."""
import ctypes

import numpy as np
import pytest

import oracle_op
from presto_b200.spi import decode_int128, encode_int128

MAX_DECIMAL = 10 ** 38 - 1
MIN_DECIMAL = -MAX_DECIMAL
U = ctypes.POINTER(ctypes.c_uint64)


def _raw(v):
    a = encode_int128([v])
    return a, a.ctypes.data_as(U)


def _out():
    a = np.zeros((1, 2), dtype=np.uint64)
    return a, a.ctypes.data_as(U)


def dec_binary(op, a, b):
    L = oracle_op.lib()
    ra, pa = _raw(a)
    rb, pb = _raw(b)
    ro, po = _out()
    st = L.oracle_dec_binary(op, pa, pb, po)
    return None if st else decode_int128(ro)[0]


def rescale(a, factor):
    L = oracle_op.lib()
    ra, pa = _raw(a)
    ro, po = _out()
    st = L.oracle_dec_rescale(pa, factor, po)
    return None if st else decode_int128(ro)[0]


def half_up_div(v, d):
    q, r = divmod(abs(v), d)
    if 2 * r >= d:
        q += 1
    return -q if v < 0 else q


RESCALE = [(10, 0, 10), (-10, 0, -10), (10, -20, 0), (15, -1, 2), (1050, -3, 1), (15, 1, 150), (-14, -1, -1), (-14, 1, -140),
           (0, 1, 0), (5, -1, 1), (10, 10, 100000000000), (150000000000000000000, -20, 2), (-140000000000000000000, -20, -1),
           (50000000000000000000, -20, 1), (150500000000000000000, -18, 151), (-140000000000000000000, -18, -140),
           (1 << 63, -18, 9), (1 << 62, -18, 5), (1 << 62, -19, 0), (MAX_DECIMAL, -1, MAX_DECIMAL // 10 + 1),
           (MIN_DECIMAL, -10, -(MAX_DECIMAL // 10000000000) - 1), (1, 37, 10 ** 37), (-1, 37, -10 ** 37), (10 ** 37, -37, 1)]


@pytest.mark.parametrize("value,factor,expected", RESCALE)
def test_rescale(value, factor, expected):
    assert rescale(value, factor) == expected


def test_rescale_overflows():
    assert rescale(1, 38) is None


ADD = [(0, 0, 0), (1, 0, 1), (1, 1, 2), (1 << 32, 0, 1 << 32), (1 << 31, 1 << 31, 1 << 32), (1 << 32, 1 << 33, (1 << 32) + (1 << 33))]


@pytest.mark.parametrize("a,b,expected", ADD)
def test_add(a, b, expected):
    assert dec_binary(0, a, b) == expected
    assert dec_binary(1, expected, b) == a


MULTIPLY = [(0, 0, 0), (1, 0, 0), (1, 1, 1), (1, -1, -1), (-1, -1, 1), (MAX_DECIMAL, 0, 0), (MAX_DECIMAL, 1, MAX_DECIMAL),
            (MIN_DECIMAL, 1, MIN_DECIMAL), (MAX_DECIMAL, -1, MIN_DECIMAL), (MIN_DECIMAL, -1, MAX_DECIMAL),
            (0xFFFFFFFFFFFFFFFF, 0xFFFFFFFFFFFFFF, 0xfffffffffffffeff00000000000001),
            (0xFFFFFF0096BFB800, 0x39003539D9A51600, 0x39003500FB00AB761CDBB17E11D00000),
            (2147483647, -2147483648, 2147483647 * -2147483648),
            (99999999999999, -1000000000000000000000000, -99999999999999000000000000000000000000),
            (12380837221737387489365741632769922889, 3, 37142511665212162468097224898309768667)]


@pytest.mark.parametrize("a,b,expected", MULTIPLY)
def test_multiply(a, b, expected):
    assert dec_binary(2, a, b) == expected
    assert dec_binary(2, b, a) == expected


def test_multiply_overflow():
    # TestUnscaledDecimal128Arithmetic.testMultiplyOverflow :200-204 (multiply throws only past 127 bits / 10^38 check)
    assert dec_binary(2, 99999999999999, -10000000000000000000000000) is None
    assert dec_binary(2, MAX_DECIMAL, 10) is None


def test_add_return_overflow_and_compare():
    L = oracle_op.lib()
    for a, b in [(2, 2), (MAX_DECIMAL, MAX_DECIMAL), (-MAX_DECIMAL, MAX_DECIMAL), (MAX_DECIMAL, -MAX_DECIMAL), (-MAX_DECIMAL, -MAX_DECIMAL)]:
        ra, pa = _raw(a)
        rb, pb = _raw(b)
        ro, po = _out()
        ov = ctypes.c_int64()
        L.oracle_dec_add_with_overflow(pa, pb, po, ctypes.byref(ov))
        s = a + b
        assert ov.value * (1 << 127) + decode_int128(ro)[0] == s or (ov.value == 0 and decode_int128(ro)[0] == s)
    # compare: zero equals negative zero (UnscaledDecimal128Arithmetic.compare :876-928)
    for a, b in [(0, 0), (1, 0), (-1, 0), (-1, 1), (MAX_DECIMAL, MIN_DECIMAL), (1 << 64, (1 << 64) - 1), (-(1 << 64), -((1 << 64) - 1))]:
        ra, pa = _raw(a)
        rb, pb = _raw(b)
        c = L.oracle_dec_compare(pa, pb)
        assert (c > 0) - (c < 0) == (a > b) - (a < b)
    negzero = np.array([[0, 1 << 63]], dtype=np.uint64)
    zero = np.zeros((1, 2), dtype=np.uint64)
    assert L.oracle_dec_compare(negzero.ctypes.data_as(U), zero.ctypes.data_as(U)) == 0
    for v, exp in [(MAX_DECIMAL, 0), (MAX_DECIMAL + 1, 1), (-(MAX_DECIMAL + 1), 1), (0, 0)]:
        rv, pv = _raw(v)
        assert L.oracle_dec_overflows(pv) == exp


def _sum_state(values):
    """LongDecimalWithOverflowState after adding `values` (DecimalSumAggregation.inputLongDecimal :141-161)."""
    L = oracle_op.lib()
    acc, pacc = _raw(0)
    overflow = 0
    for v in values:
        rv, pv = _raw(v)
        ro, po = _out()
        ov = ctypes.c_int64()
        L.oracle_dec_add_with_overflow(pacc, pv, po, ctypes.byref(ov))
        overflow += ov.value
        acc, pacc = ro, po
    return overflow, decode_int128(acc)[0]


def test_decimal_sum_state_vectors():
    T = 1 << 126
    assert _sum_state([T]) == (0, T)                                    # testOverflow
    assert _sum_state([T, T]) == (1, 0)
    assert _sum_state([-T]) == (0, -T)                                  # testUnderflow
    ov, v = _sum_state([-T, -T])
    assert ov == -1 and v == 0
    assert _sum_state([T, T, 1 << 125]) == (1, 1 << 125)                # testUnderflowAfterOverflow
    assert _sum_state([T, T, 1 << 125, -T, -T, -T]) == (0, -(1 << 125))
    # testCombineOverflow / testCombineUnderflow: combine = add the two long decimals + the overflow counters
    assert _sum_state([1 << 125, T, 1 << 125, T]) == (1, T)
    assert _sum_state([-(1 << 125), -T, -(1 << 125), -T]) == (-1, -T)
