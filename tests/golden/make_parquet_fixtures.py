#!/usr/bin/env python
"""Writes the Parquet fixtures of tests/golden/ with pyarrow's writer (run once, files committed):

    python tests/golden/make_parquet_fixtures.py

lineitem_v1.parquet  600 TPC-H-shaped rows (the columns of Q1/Q6 + orderkey), data page v1, 2 row groups, small pages,
                     dictionary encoding for the flag / discount / tax / quantity columns, PLAIN for the rest, no NULLs
nullable_v2.parquet  1 000 rows with NULLs in every column (required + optional columns, BOOLEAN, runs of equal values so the
                     hybrid streams hold both RLE and bit-packed runs), data page v2, dictionary + plain columns
Expected values travel as tests/golden/<name>.expected.npz (what pyarrow's own reader returns for the files)."""
import os

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq

HERE = os.path.dirname(os.path.abspath(__file__))


def lineitem_table(n=600, seed=5):
    rng = np.random.default_rng(seed)
    qty = rng.integers(1, 51, n).astype(np.float64)
    return pa.table({
        "orderkey": pa.array(np.repeat(np.arange(1, n // 4 + 2) * 32 // 8, 4)[:n].astype(np.int64)),
        "quantity": pa.array(qty),
        "extendedprice": pa.array(np.round(qty * rng.uniform(901, 2098.99, n), 2)),
        "discount": pa.array(rng.integers(0, 11, n) / 100.0),
        "tax": pa.array(rng.integers(0, 9, n) / 100.0),
        "returnflag": pa.array(rng.choice(["A", "N", "R"], n)),
        "linestatus": pa.array(rng.choice(["F", "O"], n)),
        "shipdate": pa.array(rng.integers(8036, 10562, n).astype(np.int32), pa.date32()),
    })


def nullable_table(n=1000, seed=6):
    rng = np.random.default_rng(seed)

    def with_nulls(values, frac, t):
        mask = rng.random(n) < frac
        return pa.array(values, t, mask=mask)
    runs = np.repeat(rng.integers(0, 7, n // 10 + 1), rng.integers(1, 40, n // 10 + 1))[:n]
    schema = pa.schema([pa.field("k", pa.int64(), nullable=False), pa.field("runs", pa.int32()), pa.field("x", pa.float64()),
                        pa.field("flag", pa.string()), pa.field("b", pa.bool_()), pa.field("allnull", pa.int64()), pa.field("big", pa.int64())])
    return pa.table({
        "k": pa.array(np.arange(n, dtype=np.int64) * 3 - 700),
        "runs": with_nulls(runs.astype(np.int32), 0.1, pa.int32()),
        "x": with_nulls(rng.normal(size=n), 0.3, pa.float64()),
        "flag": with_nulls(rng.choice(["", "a", "bb", "ccc"], n), 0.2, pa.string()),
        "b": with_nulls(rng.random(n) < 0.5, 0.15, pa.bool_()),
        "allnull": pa.array([None] * n, pa.int64()),
        "big": with_nulls(rng.integers(-2 ** 62, 2 ** 62, n), 0.05, pa.int64()),
    }, schema=schema)


def expected(table):
    out = {}
    for name in table.column_names:
        col = table.column(name).combine_chunks()
        nulls = np.asarray(col.is_null()).astype(np.uint8)
        if pa.types.is_string(col.type):
            vals = np.array([b"" if v is None else v.encode() for v in col.to_pylist()], dtype=object)
            out[name + ".values"] = np.array([v.decode() for v in vals])
        else:
            if pa.types.is_date32(col.type):
                col = col.cast(pa.int32())
            if pa.types.is_boolean(col.type):
                col = col.cast(pa.uint8())
            out[name + ".values"] = np.asarray(col.fill_null(0))
        out[name + ".nulls"] = nulls
    return out


if __name__ == "__main__":
    t = lineitem_table()
    pq.write_table(t, os.path.join(HERE, "lineitem_v1.parquet"), compression="NONE", data_page_version="1.0", data_page_size=1024,
                   row_group_size=400, use_dictionary=["quantity", "discount", "tax", "returnflag", "linestatus"], write_statistics=False, write_batch_size=50)
    np.savez(os.path.join(HERE, "lineitem_v1.expected.npz"), **expected(t))
    t = nullable_table()
    pq.write_table(t, os.path.join(HERE, "nullable_v2.parquet"), compression="NONE", data_page_version="2.0", data_page_size=512,
                   row_group_size=1000, use_dictionary=["runs", "flag", "allnull"], write_statistics=False, write_batch_size=50)
    np.savez(os.path.join(HERE, "nullable_v2.expected.npz"), **expected(t))
    for f in ("lineitem_v1.parquet", "nullable_v2.parquet"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")
