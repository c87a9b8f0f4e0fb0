// oracle/tpch_dbgen.cpp — TEST INFRASTRUCTURE (see oracle.cpp's header): CPU restatement of the TPC-H lineitem row
// generator the reference's tpch connector calls.
//
// The generator itself is NOT in the reference tree: it is the third-party artifact io.trino.tpch:tpch:1.1
// (/root/reference/pom.xml:625-627), a Java port of the TPC's dbgen, used at
// plugin/trino-tpch/src/main/java/io/trino/plugin/tpch/TpchRecordSet.java:51-60 (createTpchRecordSet ->
// table.createGenerator(scaleFactor, part, partCount)) and core/trino-main/src/test/java/io/trino/sql/gen/
// BenchmarkPageProcessor.java:120 (new LineItemGenerator(1, 1, 1)). What is restated here is the PUBLISHED algorithm of
// dbgen (TPC-H specification 2.x, clause 4.2.3, and dbgen's driver.c / build.c / rnd.c): Lehmer streams
// seed' = seed * 16807 mod (2^31 - 1), one stream per column with dbgen's seed table, every lineitem stream advanced by
// exactly O_LCNT_MAX = 7 draws per order so that any order can be generated independently.
//
// Pinned by: (1) the lineitem row counts the reference ships for the generator's output —
// plugin/trino-tpch/src/main/resources/tpch/statistics/sf0.01/lineitem.json (60175 rows) and sf1.0 (6001215 rows) —
// and the column min / max / distinct counts in the same files; (2) the reference's sf1 golden query answers
// testing/trino-product-tests/src/main/resources/sql-tests/testcases/hive_tpch/q01.result and q06.result, which the
// rows generated here reproduce to the last digit in DECIMAL arithmetic (tests/test_tpch_golden.py).
//
// Deliberately sequential and stateful (usage counters, rowFinished) like the generator it restates; the product's
// generator (presto_b200/csrc/tpch_dbgen.cuh) is a separate, random-access formulation and the tests compare the two.
#include <cstdint>
#include <cstring>

namespace {

constexpr int64_t kMultiplier = 16807;
constexpr int64_t kModulus = 2147483647;
constexpr int kLineCountMax = 7;

// one dbgen random stream: bounded ints, `perRow` draws reserved per row (order)
struct Stream {
    int64_t seed;
    int lo, hi, perRow, usage = 0;
    Stream(int64_t s, int lo_, int hi_, int perRow_ = 1) : seed(s), lo(lo_), hi(hi_), perRow(perRow_) {}
    int64_t nextRand() {
        seed = (seed * kMultiplier) % kModulus;
        usage++;
        return seed;
    }
    // dbgen's UnifInt: the double detour is part of the published behaviour
    int next() {
        nextRand();
        const int range = hi - lo + 1;
        const int v = (int)((1.0 * (double)seed / (double)kModulus) * (double)range);
        return lo + v;
    }
    void advance(int64_t count) {
        int64_t m = kMultiplier;
        while (count > 0) {
            if (count & 1) seed = (m * seed) % kModulus;
            count >>= 1;
            m = (m * m) % kModulus;
        }
    }
    void rowFinished() {
        advance(perRow - usage);
        usage = 0;
    }
    void skipRows(int64_t rows) { advance((int64_t)perRow * rows); }
};

// dbgen date indexes are days, 92001 = 1992-01-01 = epoch day 8035
constexpr int kMinGenerateDate = 92001;
constexpr int kEpochOffset = 92001 - 8035;
constexpr int kTotalDateRange = 2557;
constexpr int kCurrentDateEpoch = 9298;  // 1995-06-17 (dbgen CURRENTDATE 95168 = day 168 of 1995)
constexpr int kShipMax = 121, kReceiptMax = 30;

inline int64_t partPrice(int64_t partKey) {  // cents: P_RETAILPRICE
    return 90000 + ((partKey / 10) % 20001) + (partKey % 1000) * 100;
}

inline int64_t sparseOrderKey(int64_t orderIndex) {  // keep 3 low bits, insert 2 zero bits above them
    const int64_t low = orderIndex & 7;
    return (((orderIndex >> 3) << 2) << 3) + low;
}

}  // namespace

extern "C" {

// number of orders at a scale factor (ORDERS: 1 500 000 x sf)
int64_t oracle_tpch_order_count(double sf) { return (int64_t)(1500000.0 * sf); }

// Generates the lineitem rows of orders [first_order, first_order + n_orders) (0-based order positions; order index =
// position + 1). Any output pointer may be null. Returns the number of rows, or -(rows needed) when capacity is too small
// (nothing past capacity is written). quantity is in units, extendedprice in cents, discount / tax in hundredths.
int64_t oracle_tpch_lineitem(double sf, int64_t first_order, int64_t n_orders, int64_t capacity, int64_t* orderkey, int32_t* linenumber,
                             int64_t* quantity, int64_t* extendedprice, int64_t* discount, int64_t* tax, int32_t* shipdate, int32_t* commitdate,
                             int32_t* receiptdate, uint8_t* returnflag, uint8_t* linestatus, int64_t* partkey) {
    Stream orderDate(1066728069, kMinGenerateDate, kMinGenerateDate + kTotalDateRange - (kShipMax + kReceiptMax) - 1);
    Stream lineCount(1434868289, 1, kLineCountMax);
    Stream qty(209208115, 1, 50, kLineCountMax);
    Stream disc(554590007, 0, 10, kLineCountMax);
    Stream taxS(721958466, 0, 8, kLineCountMax);
    Stream part(1808217256, 1, (int)(int64_t)(200000.0 * sf), kLineCountMax);
    Stream suppNo(2095021727, 0, 3, kLineCountMax);
    Stream ship(1769349045, 1, kShipMax, kLineCountMax);
    Stream commit(904914315, 30, 90, kLineCountMax);
    Stream receipt(373135028, 1, kReceiptMax, kLineCountMax);
    Stream rflag(717419739, 0, 1, kLineCountMax);  // distribution "rflag": R|1, A|1
    Stream* all[] = {&orderDate, &lineCount, &qty, &disc, &taxS, &part, &suppNo, &ship, &commit, &receipt, &rflag};
    for (Stream* s : all) s->skipRows(first_order);
    int64_t row = 0;
    for (int64_t o = 0; o < n_orders; o++) {
        const int64_t orderIndex = first_order + o + 1;
        const int od = orderDate.next();
        const int lines = lineCount.next();
        for (int ln = 0; ln < lines; ln++) {
            const int q = qty.next();
            const int d = disc.next();
            const int t = taxS.next();
            const int64_t pk = part.next();
            suppNo.next();
            const int sd = ship.next() + od;
            const int cd = commit.next() + od;
            const int rd = receipt.next() + sd;
            const int sdE = sd - kEpochOffset, rdE = rd - kEpochOffset;
            uint8_t rf = 'N';
            if (rdE <= kCurrentDateEpoch) rf = rflag.next() == 0 ? 'R' : 'A';
            const uint8_t ls = sdE <= kCurrentDateEpoch ? 'F' : 'O';
            if (row < capacity) {
                if (orderkey) orderkey[row] = sparseOrderKey(orderIndex);
                if (linenumber) linenumber[row] = ln + 1;
                if (quantity) quantity[row] = q;
                if (extendedprice) extendedprice[row] = partPrice(pk) * q;
                if (discount) discount[row] = d;
                if (tax) tax[row] = t;
                if (shipdate) shipdate[row] = sdE;
                if (commitdate) commitdate[row] = cd - kEpochOffset;
                if (receiptdate) receiptdate[row] = rdE;
                if (returnflag) returnflag[row] = rf;
                if (linestatus) linestatus[row] = ls;
                if (partkey) partkey[row] = pk;
            }
            row++;
        }
        for (Stream* s : all) s->rowFinished();
    }
    return row <= capacity ? row : -row;
}

}  // extern "C"
