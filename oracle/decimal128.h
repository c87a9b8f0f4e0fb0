// oracle/decimal128.h — TEST INFRASTRUCTURE (CPU oracle). Not part of the product; only tests/, smoke() and
// bench.py's cpu_baseline / --impl reference leg may use anything under oracle/.
//
// Restatement of core/trino-spi/src/main/java/io/trino/spi/type/UnscaledDecimal128Arithmetic.java as far as
// the hot path uses it. Representation is the reference's: two 64-bit words, low word first, 127-bit magnitude,
// sign in the top bit of word 1 (:30-56).
#pragma once
#include <cstdint>
#include <stdexcept>

namespace oracle {

typedef __int128 i128;
typedef unsigned __int128 u128;

struct ArithmeticOverflow : std::runtime_error {
    ArithmeticOverflow() : std::runtime_error("Decimal overflow") {}
};

struct Dec128 {  // the 16-byte Slice
    uint64_t lo = 0, hi = 0;
    bool operator==(const Dec128& o) const { return lo == o.lo && hi == o.hi; }
};

static const uint64_t SIGN_LONG_MASK = 1ull << 63;

inline bool isNegative(const Dec128& d) { return (d.hi & SIGN_LONG_MASK) != 0; }            // :1759-1762
inline uint64_t getLong1(const Dec128& d) { return d.hi & ~SIGN_LONG_MASK; }                // getLong(decimal, 1)
inline Dec128 pack(uint64_t lo, uint64_t hi, bool negative) {                               // :1793-1800
    Dec128 r;
    r.lo = lo;
    r.hi = hi | (negative ? SIGN_LONG_MASK : 0);
    return r;
}
inline u128 magnitude(const Dec128& d) { return ((u128)getLong1(d) << 64) | d.lo; }
inline bool isZero(const Dec128& d) { return d.lo == 0 && getLong1(d) == 0; }

inline Dec128 unscaledDecimal(int64_t v) {  // :158-169
    Dec128 r;
    if (v < 0) {
        r.lo = (uint64_t)0 - (uint64_t)v;
        r.hi = SIGN_LONG_MASK;
    } else {
        r.lo = (uint64_t)v;
    }
    return r;
}

// 10^38 as magnitude
inline u128 tenTo38() { return ((u128)0x4b3b4ca85a86c47aull << 64) | 0x098a224000000000ull; }
inline bool exceedsOrEqualTenToThirtyEight(const Dec128& d) { return magnitude(d) >= tenTo38(); }  // :1815-1840
inline void throwIfOverflows(const Dec128& d) {                                                     // :1010-1015
    if (exceedsOrEqualTenToThirtyEight(d)) throw ArithmeticOverflow();
}

inline int compareAbsolute(const Dec128& a, const Dec128& b) {  // :930-950
    u128 ma = magnitude(a), mb = magnitude(b);
    return ma < mb ? -1 : ma > mb ? 1 : 0;
}

inline int compare(const Dec128& a, const Dec128& b) {  // :876-928: zero compares equal regardless of sign
    bool an = isNegative(a), bn = isNegative(b);
    if (an != bn) {
        if (isZero(a) && isZero(b)) return 0;
        return an ? -1 : 1;
    }
    int c = compareAbsolute(a, b);
    return an ? -c : c;
}

// :413-430 — ignores signs, returns the carry out of bit 126
inline int64_t addUnsignedReturnOverflow(const Dec128& l, const Dec128& r, Dec128& result, bool resultNegative) {
    uint64_t l0 = l.lo, l1 = getLong1(l), r0 = r.lo, r1 = getLong1(r);
    uint64_t z0 = l0 + r0;
    int overflow = z0 < l0 ? 1 : 0;
    uint64_t intermediate = l1 + r1 + (uint64_t)overflow;
    uint64_t z1 = intermediate & ~SIGN_LONG_MASK;
    result = pack(z0, z1, resultNegative);
    return (int64_t)(intermediate >> 63);
}

// :435-449 — assumes |left| >= |right|
inline void subtractUnsigned(const Dec128& l, const Dec128& r, Dec128& result, bool resultNegative) {
    uint64_t l0 = l.lo, l1 = getLong1(l), r0 = r.lo, r1 = getLong1(r);
    uint64_t z0 = l0 - r0;
    int underflow = l0 < z0 ? 1 : 0;
    uint64_t z1 = l1 - r1 - (uint64_t)underflow;
    result = pack(z0, z1, resultNegative);
}

// :354-379 — returns 0 / +1 (overflow) / -1 (underflow)
inline int64_t addWithOverflow(const Dec128& left, const Dec128& right, Dec128& result) {
    bool ln = isNegative(left), rn = isNegative(right);
    int64_t overflow = 0;
    if (ln == rn) {
        overflow = addUnsignedReturnOverflow(left, right, result, ln);
        if (ln) overflow = -overflow;
    } else {
        int c = compareAbsolute(left, right);
        if (c > 0) subtractUnsigned(left, right, result, ln);
        else if (c < 0) subtractUnsigned(right, left, result, !ln);
        else result = Dec128();
    }
    return overflow;
}

inline Dec128 add(const Dec128& l, const Dec128& r) {  // :340-346
    Dec128 res;
    if (addWithOverflow(l, r, res) != 0) throw ArithmeticOverflow();
    return res;
}

inline Dec128 subtract(const Dec128& left, const Dec128& right) {  // :388-411
    Dec128 res;
    bool ln = isNegative(left), rn = isNegative(right);
    if (ln != rn) {
        if (addUnsignedReturnOverflow(left, right, res, ln) != 0) throw ArithmeticOverflow();
    } else {
        int c = compareAbsolute(left, right);
        if (c > 0) subtractUnsigned(left, right, res, ln);
        else if (c < 0) subtractUnsigned(right, left, res, !ln);
        else res = Dec128();
    }
    return res;
}

// :452-650 — exact product; throws when the magnitude does not fit 127 bits
inline Dec128 multiply(const Dec128& a, const Dec128& b) {
    u128 ma = magnitude(a), mb = magnitude(b);
    bool neg = isNegative(a) != isNegative(b);
    uint64_t a0 = (uint64_t)ma, a1 = (uint64_t)(ma >> 64), b0 = (uint64_t)mb, b1 = (uint64_t)(mb >> 64);
    if (a1 != 0 && b1 != 0) throw ArithmeticOverflow();
    u128 lo = (u128)a0 * b0;
    u128 c1 = (u128)a0 * b1, c2 = (u128)a1 * b0;
    if ((c1 >> 64) != 0 || (c2 >> 64) != 0) throw ArithmeticOverflow();
    u128 mid = (u128)(uint64_t)c1 + (u128)(uint64_t)c2 + (lo >> 64);
    if ((mid >> 63) != 0) throw ArithmeticOverflow();
    return pack((uint64_t)lo, (uint64_t)mid, neg);
}

inline Dec128 powerOfTen(int n) {
    u128 v = 1;
    for (int i = 0; i < n; i++) v *= 10;
    return pack((uint64_t)v, (uint64_t)(v >> 64), false);
}

// :218-232 — rescaleFactor > 0: multiply by 10^f (overflow past 127 bits, or f >= 38 -> throws);
//            rescaleFactor < 0: divide by 10^-f rounding half up (away from zero), scaleDownRoundUp :1040-1090
inline Dec128 rescale(const Dec128& d, int rescaleFactor) {
    if (rescaleFactor == 0) return d;
    if (rescaleFactor > 0) {
        if (rescaleFactor >= 38) throw ArithmeticOverflow();
        return multiply(d, powerOfTen(rescaleFactor));
    }
    int f = -rescaleFactor;
    if (f > 38) return Dec128();
    u128 m = magnitude(d), p = magnitude(powerOfTen(f > 38 ? 38 : f));
    u128 q = m / p, rem = m % p;
    if (rem >= p - rem) q += 1;
    return pack((uint64_t)q, (uint64_t)(q >> 64), isNegative(d));  // sign kept even for a zero quotient, as the reference's pack(..., isNegative(decimal))
}

inline i128 toI128(const Dec128& d) {
    u128 m = magnitude(d);
    return isNegative(d) ? -(i128)m : (i128)m;
}
inline Dec128 fromI128(i128 v) {
    bool neg = v < 0;
    u128 m = neg ? (u128)0 - (u128)v : (u128)v;
    return pack((uint64_t)m, (uint64_t)(m >> 64) & ~SIGN_LONG_MASK, neg);
}

}  // namespace oracle
