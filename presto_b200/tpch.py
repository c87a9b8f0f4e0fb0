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
                  VariableWidthBlock, block_of, createVarcharType)

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
    st = N.lib().pgpu_gen_lineitem_host(N.LI[name], seed, first_row, rows, total_rows, out.ctypes.data)
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


# -------------------------------------------------------------------------------------------- device-resident pages
class DevicePage:
    """Columns generated directly in HBM (pgpu_gen_lineitem_device) as a pgpu_column array of device pointers."""

    def __init__(self, device: int, spec, rows: int, first_row: int = 0, total_rows: int | None = None, seed: int = SEED):
        """spec: list of (type, encoding, column name(s)) per channel:
             (type, 'plain', name)                fixed-width column
             (VARCHAR1, 'varchar1', name)         byte column + shared iota offsets
             (type, 'dict', id_name, dict_block)  dictionary ids on device, small dictionary uploaded"""
        L = N.lib()
        self.device = device
        self.rows = rows
        self._allocs = []
        self._keep = []
        total = rows if total_rows is None else total_rows
        self.cols = (N.Column * len(spec))()
        self.bytes = 0
        offsets_ptr = None

        def gen(name, extra=0):
            nbytes = (rows + extra) * np.dtype(_DT[name]).itemsize
            p = L.pgpu_device_alloc(device, max(nbytes, 16))
            if not p:
                raise MemoryError("device allocation of %d bytes failed" % nbytes)
            self._allocs.append(p)
            st = L.pgpu_gen_lineitem_device(device, N.LI[name], seed, first_row, rows, total, p)
            if st:
                raise RuntimeError("pgpu_gen_lineitem_device failed: %d" % st)
            self.bytes += nbytes
            return p

        for i, s in enumerate(spec):
            t, enc = s[0], s[1]
            col = self.cols[i]
            col.type_kind, col.precision, col.scale = t.kind, t.precision, t.scale
            col.position_count = rows
            if enc == "plain":
                col.encoding = N.ENC_PLAIN
                col.values = gen(s[2])
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
HC_DEVICE_SPEC = [(BIGINT, "plain", "orderkey"), (DOUBLE, "plain", "quantity")]
HC_SHUFFLED_DEVICE_SPEC = [(BIGINT, "plain", "orderkey_shuffled"), (DOUBLE, "plain", "quantity")]

# algorithmic bytes per row
Q6_BYTES_PER_ROW = 28
Q1_BYTES_PER_ROW = 46
HC_BYTES_PER_ROW = 16
