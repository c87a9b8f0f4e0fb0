"""Host-ingest comparison on one GPU: the same Q1 pages through pgpu_add_page (blocks, one staged copy per column) and
pgpu_add_serialized_page (the reference's wire format, one copy per page, decoded on the device). Prints one JSON line.
Auxiliary measurement (profiles/), not the bench.py contract."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import pages_serde as S  # noqa: E402  (test infrastructure: only used to PRODUCE the wire bytes a Trino worker would send)

from presto_b200 import tpch  # noqa: E402
from presto_b200.operators import GpuPipelineOperator  # noqa: E402


def main():
    rows_per_page = int(os.environ.get("PAGE_ROWS", 1 << 20))
    npages = int(os.environ.get("PAGES", 16))
    total = rows_per_page * npages
    pages = [tpch.q1_page(i * rows_per_page, rows_per_page, total) for i in range(npages)]
    wire = [S.write_raw_page(p.blocks) for p in pages]
    out = {"workload": "tpch_q1", "rows": total, "page_rows": rows_per_page, "wire_bytes_per_row": len(wire[0]) / rows_per_page}
    for name in ("blocks", "serialized"):
        best = None
        for rep in range(4):
            op = GpuPipelineOperator(tpch.q1_plan(), 0)
            t0 = time.perf_counter()
            for i, pg in enumerate(pages):
                if name == "blocks":
                    op.addInput(pg)
                else:
                    op.addSerializedInput(wire[i], rows_per_page)
            op.finish()
            res = []
            while not op.isFinished():
                p = op.getOutput()
                if p is not None:
                    res.append(p)
            dt = time.perf_counter() - t0
            op.close()
            if rep and (best is None or dt < best):
                best = dt
        out[name + "_rows_per_s"] = total / best
        out[name + "_ms"] = best * 1e3
    print(json.dumps(out))


if __name__ == "__main__":
    main()
