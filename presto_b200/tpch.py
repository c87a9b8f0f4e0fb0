"""TPC-H Q1 / Q6 / high-cardinality plans over lineitem, and the synthetic lineitem columns they run on.

Plans follow the SQL text of the reference benchmarks (testing/trino-benchmark/src/main/java/io/trino/benchmark/
SqlTpchQuery1.java, SqlTpchQuery6.java) on the tpch connector's column types
(plugin/trino-tpch/src/main/java/io/trino/plugin/tpch/TpchMetadata.java:587-600): DOUBLE quantities/prices,
DATE shipdate, VARCHAR(1) flags. Expression shapes as built in
core/trino-main/src/test/java/io/trino/sql/gen/BenchmarkPageProcessor.java:173-216.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _native as N
from .operators import Aggregation, plan_text
from .relational import and_, call, compare, constant, field
from .spi import (BIGINT, DATE, DOUBLE, DictionaryBlock, IntArrayBlock, LongArrayBlock, Page, RunLengthEncodedBlock,
                  VariableWidthBlock, block_of, createDecimalType, createVarcharType)

VARCHAR1 = createVarcharType(1)
SEED = 0x5EED0006

# Q6: channels extendedprice, discount, shipdate, quantity
Q6_INPUT = [DOUBLE, DOUBLE, DATE, DOUBLE]


def q6_plan(step: str = "single") -> str:
    f = and_(compare("ge", field(2, DATE), constant(8766, DATE)),       # shipdate >= DATE '1994-01-01'
             compare("lt", field(2, DATE), constant(9131, DATE)),       # shipdate <  DATE '1995-01-01'
             compare("ge", field(1, DOUBLE), constant(0.05, DOUBLE)),   # discount BETWEEN 0.06 - 0.01 AND 0.06 + 0.01
             compare("le", field(1, DOUBLE), constant(0.07, DOUBLE)),
             compare("lt", field(3, DOUBLE), constant(24.0, DOUBLE)))   # quantity < 24
    proj = [call("multiply", DOUBLE, field(0, DOUBLE), field(1, DOUBLE))]
    return plan_text(Q6_INPUT, f, proj, step=step, aggregations=[Aggregation("sum", DOUBLE, 0, DOUBLE)])


# Q1: channels returnflag, linestatus, quantity, extendedprice, discount, tax, shipdate
Q1_INPUT = [VARCHAR1, VARCHAR1, DOUBLE, DOUBLE, DOUBLE, DOUBLE, DATE]


def q1_plan(step: str = "single", expectedGroups: int = 10000) -> str:
    one = constant(1.0, DOUBLE)
    disc_price = call("multiply", DOUBLE, field(3, DOUBLE), call("subtract", DOUBLE, one, field(4, DOUBLE)))
    charge = call("multiply", DOUBLE, disc_price, call("add", DOUBLE, one, field(5, DOUBLE)))
    f = compare("le", field(6, DATE), constant(10471, DATE))             # shipdate <= DATE '1998-12-01' - 90 days
    proj = [field(0, VARCHAR1), field(1, VARCHAR1), field(2, DOUBLE), field(3, DOUBLE), disc_price, charge, field(4, DOUBLE)]
    aggs = [Aggregation("sum", DOUBLE, 2, DOUBLE), Aggregation("sum", DOUBLE, 3, DOUBLE), Aggregation("sum", DOUBLE, 4, DOUBLE),
            Aggregation("sum", DOUBLE, 5, DOUBLE), Aggregation("avg", DOUBLE, 2, DOUBLE), Aggregation("avg", DOUBLE, 3, DOUBLE),
            Aggregation("avg", DOUBLE, 6, DOUBLE), Aggregation("count", BIGINT)]
    return plan_text(Q1_INPUT, f, proj, step=step, groupByChannels=[0, 1], aggregations=aggs, expectedGroups=expectedGroups)


# High cardinality: GROUP BY orderkey, sum(quantity)
HC_INPUT = [BIGINT, DOUBLE]


def highcard_plan(expectedGroups: int, step: str = "single") -> str:
    return plan_text(HC_INPUT, step=step, groupByChannels=[0], aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE)],
                     expectedGroups=expectedGroups)


_DT = {"orderkey": np.int64, "quantity": np.float64, "extendedprice": np.float64, "discount": np.float64, "tax": np.float64,
       "shipdate": np.int32, "returnflag": np.uint8, "linestatus": np.uint8, "iota_offsets": np.int32,
       "discount_id": np.int32, "tax_id": np.int32, "quantity_id": np.int32, "returnflag_id": np.int32,
       "linestatus_id": np.int32, "orderkey_shuffled": np.int64}


def column_dtype(name: str):
    return _DT[name]


def host_column(name: str, first_row: int, rows: int, total_rows: int, seed: int = SEED) -> np.ndarray:
    """Rows [first_row, first_row+rows) of a synthetic lineitem column, generated on the host by the same
    counter-based generator the device uses (csrc/lineitem_gen.h)."""
    n = rows + (1 if name == "iota_offsets" else 0)
    out = np.empty(n, dtype=_DT[name])
    st = N.bench_lib().pgpu_gen_lineitem_host(N.LI[name], seed, first_row, rows, total_rows, out.ctypes.data)
    assert st == 0
    return out


def q6_page(first_row: int, rows: int, total_rows: int, seed: int = SEED) -> Page:
    c = lambda n: host_column(n, first_row, rows, total_rows, seed)
    return Page(LongArrayBlock(DOUBLE, c("extendedprice")), LongArrayBlock(DOUBLE, c("discount")),
                IntArrayBlock(DATE, c("shipdate")), LongArrayBlock(DOUBLE, c("quantity")))


def q1_page(first_row: int, rows: int, total_rows: int, seed: int = SEED) -> Page:
    c = lambda n: host_column(n, first_row, rows, total_rows, seed)
    off = c("iota_offsets")
    return Page(VariableWidthBlock(VARCHAR1, c("returnflag"), off), VariableWidthBlock(VARCHAR1, c("linestatus"), off),
                LongArrayBlock(DOUBLE, c("quantity")), LongArrayBlock(DOUBLE, c("extendedprice")),
                LongArrayBlock(DOUBLE, c("discount")), LongArrayBlock(DOUBLE, c("tax")), IntArrayBlock(DATE, c("shipdate")))


def q1_page_encoded(first_row: int, rows: int, total_rows: int, seed: int = SEED) -> Page:
    """Config 4: returnflag, linestatus, quantity, discount, tax as DictionaryBlocks (dictionaries of 3/2/50/11/9
    entries, plugin/trino-tpch/src/main/resources/tpch/statistics/sf1.0/lineitem.json); same values as q1_page."""
    c = lambda n: host_column(n, first_row, rows, total_rows, seed)
    rf = VariableWidthBlock.from_values(VARCHAR1, ["A", "N", "R"])
    ls = VariableWidthBlock.from_values(VARCHAR1, ["F", "O"])
    qty = LongArrayBlock(DOUBLE, np.arange(1, 51, dtype=np.float64))
    disc = LongArrayBlock(DOUBLE, np.arange(0, 11, dtype=np.float64) / 100.0)
    tax = LongArrayBlock(DOUBLE, np.arange(0, 9, dtype=np.float64) / 100.0)
    return Page(DictionaryBlock(rf, c("returnflag_id")), DictionaryBlock(ls, c("linestatus_id")),
                DictionaryBlock(qty, c("quantity_id")), LongArrayBlock(DOUBLE, c("extendedprice")),
                DictionaryBlock(disc, c("discount_id")), DictionaryBlock(tax, c("tax_id")), IntArrayBlock(DATE, c("shipdate")))


def highcard_page(first_row: int, rows: int, total_rows: int, shuffled: bool = False, seed: int = SEED) -> Page:
    c = lambda n: host_column(n, first_row, rows, total_rows, seed)
    return Page(LongArrayBlock(BIGINT, c("orderkey_shuffled" if shuffled else "orderkey")), LongArrayBlock(DOUBLE, c("quantity")))


# ------------------------------------------------------------------------------------------------ real TPC-H rows
# The rows the reference's tpch connector produces (csrc/tpch_dbgen.cuh), addressed by order. NATIVE = the connector's
# DOUBLE columns; DECIMAL = the DECIMAL(12,2) schema of the reference's hive_tpch golden answers (q01.result, q06.result).
DEC12_2 = createDecimalType(12, 2)
TPCH_NATIVE, TPCH_DECIMAL_12_2 = 0, 1
_TPCH_NUMERIC = ("quantity", "extendedprice", "discount", "tax")


def tpch_order_count(sf: float) -> int:
    return int(N.bench_lib().pgpu_tpch_order_count(float(sf)))


def tpch_rows(sf: float, first_order: int = 0, n_orders: int | None = None) -> int:
    n_orders = tpch_order_count(sf) - first_order if n_orders is None else n_orders
    r = ctypes.c_int64()
    if N.bench_lib().pgpu_tpch_lineitem_rows(float(sf), first_order, n_orders, ctypes.byref(r)):
        raise ValueError("order range [%d, %d) is outside scale factor %s" % (first_order, first_order + n_orders, sf))
    return r.value


def _tpch_dtype(name: str, decimal: bool):
    return np.int64 if decimal and name in _TPCH_NUMERIC else _DT[name]


def tpch_host_column(name: str, sf: float, first_order: int, n_orders: int, decimal: bool = False, rows: int | None = None) -> np.ndarray:
    rows = tpch_rows(sf, first_order, n_orders) if rows is None else rows
    if name == "iota_offsets":
        return np.arange(rows + 1, dtype=np.int32)
    out = np.empty(rows, dtype=_tpch_dtype(name, decimal))
    fmt = TPCH_DECIMAL_12_2 if decimal and name in _TPCH_NUMERIC else TPCH_NATIVE
    st = N.bench_lib().pgpu_tpch_lineitem_host(float(sf), first_order, n_orders, N.LI[name], fmt, rows, out.ctypes.data, None)
    assert st == 0, st
    return out


def tpch_q1_page(sf: float, first_order: int, n_orders: int, decimal: bool = False) -> Page:
    rows = tpch_rows(sf, first_order, n_orders)
    c = lambda n: tpch_host_column(n, sf, first_order, n_orders, decimal, rows)
    t = DEC12_2 if decimal else DOUBLE
    off = c("iota_offsets")
    return Page(VariableWidthBlock(VARCHAR1, c("returnflag"), off), VariableWidthBlock(VARCHAR1, c("linestatus"), off),
                LongArrayBlock(t, c("quantity")), LongArrayBlock(t, c("extendedprice")), LongArrayBlock(t, c("discount")),
                LongArrayBlock(t, c("tax")), IntArrayBlock(DATE, c("shipdate")))


def tpch_q6_page(sf: float, first_order: int, n_orders: int, decimal: bool = False) -> Page:
    rows = tpch_rows(sf, first_order, n_orders)
    c = lambda n: tpch_host_column(n, sf, first_order, n_orders, decimal, rows)
    t = DEC12_2 if decimal else DOUBLE
    return Page(LongArrayBlock(t, c("extendedprice")), LongArrayBlock(t, c("discount")), IntArrayBlock(DATE, c("shipdate")),
                LongArrayBlock(t, c("quantity")))


def tpch_highcard_page(sf: float, first_order: int, n_orders: int) -> Page:
    rows = tpch_rows(sf, first_order, n_orders)
    c = lambda n: tpch_host_column(n, sf, first_order, n_orders, False, rows)
    return Page(LongArrayBlock(BIGINT, c("orderkey")), LongArrayBlock(DOUBLE, c("quantity")))


def q1_decimal_plan(step: str = "single", expectedGroups: int = 10000) -> str:
    """hive_tpch/q01.sql on DECIMAL(12,2) columns, typed as the reference's analyzer types it: the integer literal 1 is
    coerced to decimal(10,0), so 1 - l_discount and 1 + l_tax are decimal(13,2), the products decimal(25,4) and
    decimal(38,6) (main/type/DecimalOperators.java:84-85,261-262); sum -> decimal(38,s), avg -> the input type."""
    d10, d13, d25, d38_6 = createDecimalType(10, 0), createDecimalType(13, 2), createDecimalType(25, 4), createDecimalType(38, 6)
    one = constant(1, d10)
    disc_price = call("multiply", d25, field(3, DEC12_2), call("subtract", d13, one, field(4, DEC12_2)))
    charge = call("multiply", d38_6, disc_price, call("add", d13, one, field(5, DEC12_2)))
    f = compare("le", field(6, DATE), constant(10471, DATE))
    proj = [field(0, VARCHAR1), field(1, VARCHAR1), field(2, DEC12_2), field(3, DEC12_2), disc_price, charge, field(4, DEC12_2)]
    s2, s4 = createDecimalType(38, 2), createDecimalType(38, 4)
    aggs = [Aggregation("sum", s2, 2, DEC12_2), Aggregation("sum", s2, 3, DEC12_2), Aggregation("sum", s4, 4, d25),
            Aggregation("sum", d38_6, 5, d38_6), Aggregation("avg", DEC12_2, 2, DEC12_2), Aggregation("avg", DEC12_2, 3, DEC12_2),
            Aggregation("avg", DEC12_2, 6, DEC12_2), Aggregation("count", BIGINT)]
    inputs = [VARCHAR1, VARCHAR1, DEC12_2, DEC12_2, DEC12_2, DEC12_2, DATE]
    return plan_text(inputs, f, proj, step=step, groupByChannels=[0, 1], aggregations=aggs, expectedGroups=expectedGroups)


def q6_decimal_plan(step: str = "single") -> str:
    """hive_tpch/q06.sql on DECIMAL(12,2) columns; literals as the optimizer leaves them after folding and coercion to the
    column type (0.05, 0.07, 24 as decimal(12,2)); extendedprice * discount is decimal(24,4), its sum decimal(38,4)."""
    d24 = createDecimalType(24, 4)
    f = and_(compare("ge", field(2, DATE), constant(8766, DATE)), compare("lt", field(2, DATE), constant(9131, DATE)),
             compare("ge", field(1, DEC12_2), constant(5, DEC12_2)), compare("le", field(1, DEC12_2), constant(7, DEC12_2)),
             compare("lt", field(3, DEC12_2), constant(2400, DEC12_2)))
    proj = [call("multiply", d24, field(0, DEC12_2), field(1, DEC12_2))]
    return plan_text([DEC12_2, DEC12_2, DATE, DEC12_2], f, proj, step=step, aggregations=[Aggregation("sum", createDecimalType(38, 4), 0, d24)])


# -------------------------------------------------------------------------------------------- device-resident pages
class DevicePage:
    """Columns generated directly in HBM (pgpu_gen_lineitem_device) as a pgpu_column array of device pointers."""

    def __init__(self, device: int, spec, rows: int | None = None, first_row: int = 0, total_rows: int | None = None, seed: int = SEED, dbgen=None):
        """spec: list of (type, encoding, column name(s)) per channel:
             (type, 'plain', name)                fixed-width column
             (VARCHAR1, 'varchar1', name)         byte column + shared iota offsets
             (type, 'dict', id_name, dict_block)  dictionary ids on device, small dictionary uploaded
           dbgen = (scale factor, first order, orders): the connector's real TPC-H rows of that order range
           (pgpu_tpch_lineitem_device; `rows` is then derived) instead of the synthetic columns; DECIMAL(12,2)-typed
           channels get the unscaled int64 form."""
        L = N.lib()
        self.device = device
        if dbgen is not None:
            rows = tpch_rows(*dbgen)
        self.rows = rows
        self._allocs = []
        self._keep = []
        total = rows if total_rows is None else total_rows
        self.cols = (N.Column * len(spec))()
        self.bytes = 0
        offsets_ptr = None

        def gen(name, extra=0, fmt=TPCH_NATIVE):
            nbytes = (rows + extra) * np.dtype(_DT[name]).itemsize
            p = L.pgpu_device_alloc(device, max(nbytes, 16))
            if not p:
                raise MemoryError("device allocation of %d bytes failed" % nbytes)
            self._allocs.append(p)
            if dbgen is not None and name != "iota_offsets":
                sf, first_order, n_orders = dbgen
                st = L.pgpu_tpch_lineitem_device(device, float(sf), first_order, n_orders, N.LI[name], fmt, rows, p, None)
            else:
                st = L.pgpu_gen_lineitem_device(device, N.LI[name], seed, first_row, rows, total, p)
            if st:
                raise RuntimeError("lineitem generation on the device failed: %d" % st)
            self.bytes += nbytes
            return p

        for i, s in enumerate(spec):
            t, enc = s[0], s[1]
            col = self.cols[i]
            col.type_kind, col.precision, col.scale = t.kind, t.precision, t.scale
            col.position_count = rows
            if enc == "plain":
                col.encoding = N.ENC_PLAIN
                col.values = gen(s[2], fmt=TPCH_DECIMAL_12_2 if t == DEC12_2 else TPCH_NATIVE)
            elif enc == "varchar1":
                col.encoding = N.ENC_PLAIN
                col.values = gen(s[2])
                offsets_ptr = gen("iota_offsets", 1)  # every VariableWidthBlock owns its offsets array
                col.offsets = offsets_ptr
            elif enc == "dict":
                col.encoding = N.ENC_DICTIONARY
                col.ids = gen(s[2])
                dblock = s[3]
                d = N.Column()
                keep = []
                dblock._fill(d, keep)
                # upload the (tiny) dictionary
                def up(ptr, nbytes):
                    if not ptr:
                        return None
                    p = L.pgpu_device_alloc(device, max(nbytes, 16))
                    self._allocs.append(p)
                    if L.pgpu_memcpy_h2d(device, p, ptr, nbytes):
                        raise RuntimeError("dictionary upload failed")
                    return p
                n = dblock.position_count
                if t.is_var_width:
                    d.values = up(d.values, int(dblock.offsets[-1]) or 1)
                    d.offsets = up(d.offsets, (n + 1) * 4)
                else:
                    width = 16 if t.is_long_decimal else (1 if t.kind == 0 else 4 if t.kind in (2, 3) else 8)
                    d.values = up(d.values, n * width)
                if d.nulls:
                    d.nulls = up(d.nulls, n)
                self._keep.append(keep)
                self._keep.append(d)
                col.dictionary = ctypes.pointer(d)
            else:
                raise ValueError(enc)

    def free(self):
        L = N.lib()
        for p in self._allocs:
            L.pgpu_device_free(self.device, p)
        self._allocs = []

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


Q6_DEVICE_SPEC = [(DOUBLE, "plain", "extendedprice"), (DOUBLE, "plain", "discount"), (DATE, "plain", "shipdate"),
                  (DOUBLE, "plain", "quantity")]
Q1_DEVICE_SPEC = [(VARCHAR1, "varchar1", "returnflag"), (VARCHAR1, "varchar1", "linestatus"), (DOUBLE, "plain", "quantity"),
                  (DOUBLE, "plain", "extendedprice"), (DOUBLE, "plain", "discount"), (DOUBLE, "plain", "tax"),
                  (DATE, "plain", "shipdate")]
Q6_DECIMAL_DEVICE_SPEC = [(DEC12_2, "plain", "extendedprice"), (DEC12_2, "plain", "discount"), (DATE, "plain", "shipdate"),
                          (DEC12_2, "plain", "quantity")]
Q1_DECIMAL_DEVICE_SPEC = [(VARCHAR1, "varchar1", "returnflag"), (VARCHAR1, "varchar1", "linestatus"), (DEC12_2, "plain", "quantity"),
                          (DEC12_2, "plain", "extendedprice"), (DEC12_2, "plain", "discount"), (DEC12_2, "plain", "tax"),
                          (DATE, "plain", "shipdate")]
HC_DEVICE_SPEC = [(BIGINT, "plain", "orderkey"), (DOUBLE, "plain", "quantity")]
HC_SHUFFLED_DEVICE_SPEC = [(BIGINT, "plain", "orderkey_shuffled"), (DOUBLE, "plain", "quantity")]

# algorithmic bytes per row
Q6_BYTES_PER_ROW = 28
Q1_BYTES_PER_ROW = 46
HC_BYTES_PER_ROW = 16
