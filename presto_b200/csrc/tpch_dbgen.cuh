// tpch_dbgen.cuh — TPC-H lineitem rows as the reference's tpch connector produces them, in a random-access form that one
// GPU thread per ORDER can evaluate (host code compiles the same functions).
//
// The reference gets its rows from the un-vendored io.trino.tpch:tpch:1.1 (a Java port of the TPC's dbgen), called at
// plugin/trino-tpch/src/main/java/io/trino/plugin/tpch/TpchRecordSet.java:51-60; column types and the long -> double
// conversions are the connector's (TpchMetadata.java:587-600, TpchRecordSet.java:158-196: quantity as is,
// extendedprice / discount / tax = hundredths / 100.0, dates as epoch days). dbgen's published algorithm: every column is
// a Lehmer stream seed' = 16807 * seed mod (2^31 - 1) with its own start seed, and every lineitem stream moves on by exactly
// 7 draws (the maximum line count) per order whatever the order's real line count. Hence the state of every stream at
// the start of order o is seed0 * 16807^(7 o) — two modular powers per order (one for the per-order streams, one for the
// per-line streams), shared by all columns — and an order's <= 7 lines are produced by one thread with no communication.
// The modulus is a Mersenne prime, so a 62-bit product is reduced with two shift-and-add folds instead of a division.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define DBGEN_HD __host__ __device__ __forceinline__
#else
#define DBGEN_HD inline
#endif

namespace dbgen {

constexpr uint64_t kModulus = 2147483647ull;  // 2^31 - 1
constexpr uint64_t kMultiplier = 16807ull;
constexpr int kMaxLines = 7;
// dbgen's seed table (driver.c): O_ODATE, O_LCNT, L_QTY, L_DCNT, L_TAX, L_PKEY, L_SKEY, L_SDTE, L_CDTE, L_RDTE, L_RFLG
constexpr uint32_t kSeedOrderDate = 1066728069u, kSeedLineCount = 1434868289u, kSeedQuantity = 209208115u, kSeedDiscount = 554590007u,
                   kSeedTax = 721958466u, kSeedPartKey = 1808217256u, kSeedShipDate = 1769349045u, kSeedCommitDate = 904914315u,
                   kSeedReceiptDate = 373135028u, kSeedReturnFlag = 717419739u;
constexpr int kOrderDateMin = 8035;                          // 1992-01-01 as an epoch day (dbgen STARTDATE 92001)
constexpr int kOrderDateSpan = 2557 - (121 + 30) - 1 + 1;    // TOTDATE - (L_SDTE_MAX + L_RDTE_MAX) - 1, inclusive range
constexpr int kCurrentDate = 9298;                           // 1995-06-17 (dbgen CURRENTDATE 95168)

DBGEN_HD uint32_t mulmod(uint64_t a, uint64_t b) {
    uint64_t x = a * b;                     // < 2^62
    x = (x & kModulus) + (x >> 31);         // < 2^32
    x = (x & kModulus) + (x >> 31);         // <= 2^31
    return (uint32_t)(x >= kModulus ? x - kModulus : x);
}

DBGEN_HD uint32_t powmod(uint64_t n) {  // 16807^n
    uint64_t m = kMultiplier, r = 1;
    while (n) {
        if (n & 1) r = mulmod(r, m);
        m = mulmod(m, m);
        n >>= 1;
    }
    return (uint32_t)r;
}

// the stream's next state and its value in [lo, lo + range): dbgen takes the detour through double, so do we
DBGEN_HD int draw(uint32_t& seed, int lo, int range) {
    seed = mulmod(seed, kMultiplier);
    return lo + (int)(((double)seed / (double)kModulus) * (double)range);
}

DBGEN_HD int64_t orderCount(double sf) { return (int64_t)(1500000.0 * sf); }

DBGEN_HD int64_t sparseOrderKey(int64_t orderIndex) {  // 8 consecutive keys, then a gap of 24
    return ((orderIndex >> 3) << 5) + (orderIndex & 7);
}

DBGEN_HD int64_t retailPriceCents(int64_t partKey) { return 90000 + ((partKey / 10) % 20001) + (partKey % 1000) * 100; }

struct Order {
    uint32_t g7;       // 16807^(7 * position): multiplies every per-line stream's start seed
    int orderDate;     // epoch day
    int lines;         // 1..7
};

// order position (0-based; the order's index in dbgen is position + 1)
DBGEN_HD Order orderAt(int64_t pos) {
    Order o;
    const uint32_t g1 = powmod((uint64_t)pos);
    uint32_t s = mulmod(kSeedOrderDate, g1);
    o.orderDate = draw(s, kOrderDateMin, kOrderDateSpan);
    s = mulmod(kSeedLineCount, g1);
    o.lines = draw(s, 1, kMaxLines);
    uint32_t g2 = mulmod(g1, g1), g4 = mulmod(g2, g2);
    o.g7 = mulmod(mulmod(g4, g2), g1);
    return o;
}

// line count alone (row offsets need only this)
DBGEN_HD int lineCountAt(int64_t pos) {
    uint32_t s = mulmod(kSeedLineCount, powmod((uint64_t)pos));
    return draw(s, 1, kMaxLines);
}

struct Line {
    int quantity;            // 1..50
    int discount, tax;       // hundredths
    int64_t partKey;
    int64_t extendedPrice;   // cents
    int shipDate, commitDate, receiptDate;  // epoch days
    char returnFlag, lineStatus;
};

// walks the lines of one order in line-number order
struct LineCursor {
    uint32_t sQty, sDisc, sTax, sPart, sShip, sCommit, sReceipt, sFlag;
    int orderDate, partRange;
    DBGEN_HD LineCursor(const Order& o, double sf) {
        sQty = mulmod(kSeedQuantity, o.g7);
        sDisc = mulmod(kSeedDiscount, o.g7);
        sTax = mulmod(kSeedTax, o.g7);
        sPart = mulmod(kSeedPartKey, o.g7);
        sShip = mulmod(kSeedShipDate, o.g7);
        sCommit = mulmod(kSeedCommitDate, o.g7);
        sReceipt = mulmod(kSeedReceiptDate, o.g7);
        sFlag = mulmod(kSeedReturnFlag, o.g7);
        orderDate = o.orderDate;
        partRange = (int)(int64_t)(200000.0 * sf);
    }
    DBGEN_HD Line next() {
        Line l;
        l.quantity = draw(sQty, 1, 50);
        l.discount = draw(sDisc, 0, 11);
        l.tax = draw(sTax, 0, 9);
        l.partKey = draw(sPart, 1, partRange);
        l.extendedPrice = retailPriceCents(l.partKey) * l.quantity;
        l.shipDate = orderDate + draw(sShip, 1, 121);
        l.commitDate = orderDate + draw(sCommit, 30, 61);
        l.receiptDate = l.shipDate + draw(sReceipt, 1, 30);
        // the return-flag stream is only consumed by lines already received on CURRENTDATE ("rflag": R|1, A|1)
        l.returnFlag = 'N';
        if (l.receiptDate <= kCurrentDate) l.returnFlag = draw(sFlag, 0, 2) == 0 ? 'R' : 'A';
        l.lineStatus = l.shipDate <= kCurrentDate ? 'F' : 'O';
        return l;
    }
};

}  // namespace dbgen
