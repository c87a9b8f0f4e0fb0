"""Host-ingest comparison on one GPU: Q1 over the same rows through pgpu_add_page (plain blocks) and through
pgpu_add_parquet_row_group (Parquet row groups written by pyarrow, COMPRESSION=NONE|SNAPPY|LZ4: dictionary-encoded flag / quantity / discount /
tax chunks, PLAIN extendedprice / shipdate). Prints one JSON line with rows/s of both and the file bytes per row.
Auxiliary measurement (profiles/), not the bench.py contract.   PAGE_ROWS / PAGES set the size."""
import io
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import pyarrow as pa  # noqa: E402
import pyarrow.parquet as pq  # noqa: E402

from presto_b200 import parquet, tpch  # noqa: E402
from presto_b200.operators import GpuPipelineOperator  # noqa: E402

COLUMNS = ["returnflag", "linestatus", "quantity", "extendedprice", "discount", "tax", "shipdate"]


def main():
    rows_per_group = int(os.environ.get("PAGE_ROWS", 1 << 20))
    ngroups = int(os.environ.get("PAGES", 16))
    total = rows_per_group * ngroups
    pages = [tpch.q1_page(i * rows_per_group, rows_per_group, total) for i in range(ngroups)]
    files = []
    for pg in pages:
        b = pg.blocks
        t = pa.table({"returnflag": pa.array([chr(x) for x in b[0].data[:rows_per_group]]), "linestatus": pa.array([chr(x) for x in b[1].data[:rows_per_group]]),
                      "quantity": pa.array(b[2].values.view("f8")), "extendedprice": pa.array(b[3].values.view("f8")), "discount": pa.array(b[4].values.view("f8")),
                      "tax": pa.array(b[5].values.view("f8")), "shipdate": pa.array(b[6].values, pa.date32())})
        buf = io.BytesIO()
        pq.write_table(t, buf, compression=os.environ.get("COMPRESSION", "NONE"), use_dictionary=["returnflag", "linestatus", "quantity", "discount", "tax"], row_group_size=rows_per_group,
                       data_page_size=1 << 20, write_statistics=False)
        files.append(list(parquet.row_groups(buf.getvalue())))
    file_bytes = sum(len(c[0]) for f in files for _, cols in f for c in cols.values())
    out = {"workload": "tpch_q1", "rows": total, "row_group_rows": rows_per_group, "parquet_bytes_per_row": file_bytes / total,
           "compression": os.environ.get("COMPRESSION", "NONE")}
    results = {}
    for name in ("blocks", "parquet"):
        best = None
        for rep in range(4):
            op = GpuPipelineOperator(tpch.q1_plan(), 0)
            t0 = time.perf_counter()
            for i in range(ngroups):
                if name == "blocks":
                    op.addInput(pages[i])
                else:
                    for rows, cols in files[i]:
                        op.addParquetRowGroup([cols[c] for c in COLUMNS], rows)
            op.finish()
            res = []
            while not op.isFinished():
                p = op.getOutput()
                if p is not None:
                    res.append(p)
            dt = time.perf_counter() - t0
            results[name] = sorted(r for p in res for r in p.to_rows())
            op.close()
            if rep and (best is None or dt < best):
                best = dt
        out[name + "_rows_per_s"] = total / best
        out[name + "_ms"] = best * 1e3
    out["same_groups_and_counts"] = [r[:2] + r[-1:] for r in results["blocks"]] == [r[:2] + r[-1:] for r in results["parquet"]]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
