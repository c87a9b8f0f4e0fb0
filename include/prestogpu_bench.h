/*
 * prestogpu_bench.h — bench / test plumbing that is NOT part of the drop-in boundary: synthetic and spec-conformant TPC-H
 * lineitem generators (host + device), synchronous copies for callers without torch, an L2 flush. Exported by
 * libprestogpu_bench.so (csrc/gen.cu), never by libprestogpu.so: `nm -D libprestogpu.so` shows none of these.
 */
#ifndef PRESTOGPU_BENCH_H
#define PRESTOGPU_BENCH_H

#include "prestogpu.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Synchronous copies between host and device memory (test / bench plumbing for callers without torch). */
pgpu_status pgpu_memcpy_h2d(int32_t device, void* dev_dst, const void* host_src, size_t bytes);
pgpu_status pgpu_memcpy_d2h(int32_t device, void* host_dst, const void* dev_src, size_t bytes);

/* Synthetic TPC-H lineitem columns generated directly in HBM (SURVEY.md §8(d) distributions; the
 * reference's generator io.trino.tpch:tpch:1.1 is not in its tree). Column ids: see PGPU_LI_* below.
 * Row r of a column depends only on (seed, r): pgpu_gen_lineitem_host produces the identical bytes on the CPU. */
#define PGPU_LI_ORDERKEY 0      /* int64  */
#define PGPU_LI_QUANTITY 1      /* double */
#define PGPU_LI_EXTENDEDPRICE 2 /* double */
#define PGPU_LI_DISCOUNT 3      /* double */
#define PGPU_LI_TAX 4           /* double */
#define PGPU_LI_SHIPDATE 5      /* int32 days since epoch */
#define PGPU_LI_RETURNFLAG 6    /* 1 byte per row (varchar(1) payload) */
#define PGPU_LI_LINESTATUS 7    /* 1 byte per row */
#define PGPU_LI_IOTA_OFFSETS 8  /* int32 offsets 0..rows (rows+1 entries) for varchar(1) blocks */
#define PGPU_LI_DISCOUNT_ID 9   /* int32 dictionary id of discount (0..10) */
#define PGPU_LI_TAX_ID 10       /* int32 dictionary id of tax (0..8) */
#define PGPU_LI_QUANTITY_ID 11  /* int32 dictionary id of quantity (0..49) */
#define PGPU_LI_RETURNFLAG_ID 12/* int32 dictionary id of returnflag (A,N,R) */
#define PGPU_LI_LINESTATUS_ID 13/* int32 dictionary id of linestatus (F,O) */
#define PGPU_LI_ORDERKEY_SHUFFLED 14 /* int64, uniform random over the same key domain */
pgpu_status pgpu_gen_lineitem_device(int32_t device, int32_t column, uint64_t seed, int64_t first_row, int64_t rows,
                                     int64_t total_rows, void* dev_out);
pgpu_status pgpu_gen_lineitem_host(int32_t column, uint64_t seed, int64_t first_row, int64_t rows,
                                   int64_t total_rows, void* host_out);

/* TPC-H lineitem rows exactly as the reference's tpch connector produces them (plugin/trino-tpch/.../TpchRecordSet.java:51-60
 * calling the un-vendored io.trino.tpch:tpch:1.1 generator = the TPC's dbgen): same rows, same order, at any scale factor.
 * Rows are addressed by ORDER: orders [first_order, first_order + n_orders) of pgpu_tpch_order_count(sf) = 1 500 000 x sf, each
 * with 1..7 lineitems; pgpu_tpch_lineitem_rows gives the row count of a range (size the buffers with it). Columns are the
 * PGPU_LI_* ids above (not IOTA_OFFSETS / ORDERKEY_SHUFFLED). Format NATIVE = the connector's types (TpchMetadata.java:587-600:
 * DOUBLE quantity / extendedprice / discount / tax, int32 epoch days, one byte per flag); DECIMAL_12_2 = unscaled int64 of the
 * DECIMAL(12,2) columns the reference's hive_tpch golden answers (q01.result, q06.result) were computed on.
 * The generator is pinned by those answers and by the reference's table statistics (tests/test_tpch_golden.py). */
#define PGPU_TPCH_NATIVE 0
#define PGPU_TPCH_DECIMAL_12_2 1
int64_t pgpu_tpch_order_count(double sf);
pgpu_status pgpu_tpch_lineitem_rows(double sf, int64_t first_order, int64_t n_orders, int64_t* rows);
pgpu_status pgpu_tpch_lineitem_device(int32_t device, double sf, int64_t first_order, int64_t n_orders, int32_t column, int32_t format,
                                      int64_t capacity_rows, void* dev_out, int64_t* rows);
pgpu_status pgpu_tpch_lineitem_host(double sf, int64_t first_order, int64_t n_orders, int32_t column, int32_t format,
                                    int64_t capacity_rows, void* host_out, int64_t* rows);
/* The product's staging copy alone (csrc/hostcopy.cpp: worker pool, non-temporal stores) from `src` (pageable) to `dst` (pinned):
 * lets bench.py measure the host-side ceiling of the ingest path with no GPU involved. */
pgpu_status pgpu_bench_host_copy(void* dst, const void* src, size_t bytes);
/* Write `bytes` of device memory (L2 flush between timed iterations). */
pgpu_status pgpu_flush_l2(int32_t device, void* dev_scratch, size_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* PRESTOGPU_BENCH_H */
