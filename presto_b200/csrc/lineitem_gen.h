// lineitem_gen.h — synthetic TPC-H lineitem columns as a pure function of (seed, row).
//
// The reference's generator (io.trino.tpch:tpch:1.1, call site plugin/trino-tpch/.../TpchRecordSet.java:51-60) is
// not in its tree, so benchmarks and parity tests use this counter-based generator with the TPC-H value
// distributions of plugin/trino-tpch/src/main/resources/tpch/statistics/sf1.0/lineitem.json:
//   orders: 7 consecutive orders hold 1..7 lines (28 rows, mean 4.0 lines/order as in TPC-H); order keys are sparse
//           exactly like dbgen's (8 used out of every 32): key = (o / 8) * 32 + o % 8 + 1
//   orderdate  U[8035, 10440] per order; shipdate = orderdate + U[1,121]  -> [8036, 10561]
//   quantity   U{1..50}; discount {0.00..0.10}; tax {0.00..0.08}
//   extendedprice = quantity * partprice, partprice U[901.00, 2098.99] (cents exact)
//   linestatus 'O' if shipdate > 1995-06-17 (9298) else 'F'
//   returnflag 'N' if receiptdate (= shipdate + U[1,30]) > 9298 else 'R' or 'A' with equal probability
// The same code compiles for the host (oracle inputs, CPU baseline) and the device (bench inputs), so both sides see
// identical bytes.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define LI_HD __host__ __device__ __forceinline__
#else
#define LI_HD inline
#endif

namespace ligen {

LI_HD uint64_t mix(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
    x ^= x >> 27; x *= 0x94d049bb133111ebull;
    x ^= x >> 31;
    return x;
}
LI_HD uint64_t rnd(uint64_t seed, uint64_t idx, uint64_t stream) { return mix(seed ^ mix(idx * 0x9e3779b97f4a7c15ull + stream * 0xd1b54a32d192ed03ull + 1)); }
LI_HD uint32_t uniform(uint64_t r, uint32_t n) { return (uint32_t)(((r >> 32) * (uint64_t)n) >> 32); }  // [0, n)

struct Row {
    int64_t orderkey;
    int32_t quantity;      // 1..50
    int64_t price_cents;   // extendedprice in cents
    int32_t discount;      // 0..10 (hundredths)
    int32_t tax;           // 0..8
    int32_t shipdate;
    char returnflag, linestatus;
};

LI_HD int64_t order_of_row(int64_t row) {
    // 28-row blocks of 7 orders with 1..7 lines: prefix sums 0,1,3,6,10,15,21
    int64_t block = row / 28;
    int pos = (int)(row % 28);
    int j = pos >= 21 ? 6 : pos >= 15 ? 5 : pos >= 10 ? 4 : pos >= 6 ? 3 : pos >= 3 ? 2 : pos >= 1 ? 1 : 0;
    return block * 7 + j;
}
LI_HD int64_t key_of_order(int64_t o) { return (o / 8) * 32 + o % 8 + 1; }

LI_HD Row row(uint64_t seed, int64_t r) {
    Row x;
    int64_t o = order_of_row(r);
    x.orderkey = key_of_order(o);
    int32_t orderdate = 8035 + (int32_t)uniform(rnd(seed, (uint64_t)o, 1), 10440 - 8035 + 1);
    uint64_t a = rnd(seed, (uint64_t)r, 2), b = rnd(seed, (uint64_t)r, 3);
    x.quantity = 1 + (int32_t)uniform(a, 50);
    x.discount = (int32_t)uniform(a << 20, 11);
    x.tax = (int32_t)uniform(a << 40, 9);
    int64_t part_cents = 90100 + (int64_t)uniform(b, 209899 - 90100 + 1);
    x.price_cents = part_cents * x.quantity;
    x.shipdate = orderdate + 1 + (int32_t)uniform(b << 24, 121);
    int32_t receipt = x.shipdate + 1 + (int32_t)uniform(b << 44, 30);
    x.linestatus = x.shipdate > 9298 ? 'O' : 'F';
    x.returnflag = receipt > 9298 ? 'N' : ((rnd(seed, (uint64_t)r, 4) >> 63) ? 'R' : 'A');
    return x;
}

// Writes element `r` of column `column` (PGPU_LI_* ids of include/prestogpu.h) to out[r - first].
LI_HD void store(int column, uint64_t seed, int64_t r, int64_t first, int64_t total_rows, void* out) {
    int64_t i = r - first;
    if (column == 8) {  // iota offsets: rows + 1 entries
        ((int32_t*)out)[i] = (int32_t)i;
        return;
    }
    if (column == 14) {  // shuffled order keys: uniform over the same key domain
        int64_t norders = (total_rows + 27) / 28 * 7;
        uint64_t h = rnd(seed, (uint64_t)r, 5);
        int64_t o = (int64_t)(((unsigned __int128)h * (uint64_t)norders) >> 64);
        ((int64_t*)out)[i] = key_of_order(o);
        return;
    }
    Row x = row(seed, r);
    switch (column) {
        case 0: ((int64_t*)out)[i] = x.orderkey; break;
        case 1: ((double*)out)[i] = (double)x.quantity; break;
        case 2: ((double*)out)[i] = (double)x.price_cents / 100.0; break;
        case 3: ((double*)out)[i] = (double)x.discount / 100.0; break;
        case 4: ((double*)out)[i] = (double)x.tax / 100.0; break;
        case 5: ((int32_t*)out)[i] = x.shipdate; break;
        case 6: ((char*)out)[i] = x.returnflag; break;
        case 7: ((char*)out)[i] = x.linestatus; break;
        case 9: ((int32_t*)out)[i] = x.discount; break;
        case 10: ((int32_t*)out)[i] = x.tax; break;
        case 11: ((int32_t*)out)[i] = x.quantity - 1; break;
        case 12: ((int32_t*)out)[i] = x.returnflag == 'A' ? 0 : x.returnflag == 'N' ? 1 : 2; break;
        case 13: ((int32_t*)out)[i] = x.linestatus == 'F' ? 0 : 1; break;
        default: break;
    }
}

}  // namespace ligen
