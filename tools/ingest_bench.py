#!/usr/bin/env python
"""End-to-end ingest at the reference's page shape: Q1 over host pages of --page-rows rows in ordinary
pageable numpy memory through pgpu_add_page(flags=0) — the call the JNI shim makes for Java heap arrays — and through
pgpu_add_serialized_page, next to a plain pinned H2D ceiling measured in the same process.

  python tools/ingest_bench.py [--rows N] [--page-rows 1000000,65536,4096] [--steps K] [--workload q1|q6|highcard|q1dict]
"""
import argparse
import ctypes
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=128_000_000)
    ap.add_argument("--page-rows", default="1000000,65536,4096")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--workload", default="q1")
    ap.add_argument("--serialized", action="store_true")
    args = ap.parse_args()
    import numpy as np
    import torch
    import bench as B
    from presto_b200 import _native as N
    from presto_b200 import tpch
    from presto_b200.operators import GpuPipelineOperator
    L = N.lib()
    device = 0
    torch.cuda.set_device(device)
    wl = B.workload_def(args.workload, args.rows)
    wl["name"] = args.workload
    out = {"workload": args.workload, "rows": args.rows, "host_cores": os.cpu_count(), "copy_threads_env": os.environ.get("PGPU_COPY_THREADS"),
           "stage_mb": os.environ.get("PGPU_STAGE_MB", "64")}
    out["h2d_ceiling_gbs"] = B.h2d_ceiling(device)
    dp = tpch.DevicePage(device, wl["spec"], args.rows)
    host = B.HostTable(device, wl, dp, args.rows)
    out["bytes_per_row"] = host.bytes_per_row
    op = GpuPipelineOperator(wl["plan"]("single"), device)
    small = args.workload in ("q1", "q1dict", "q6")
    for pr in [int(x) for x in args.page_rows.split(",")]:
        pages = host.pages(pr)
        r = B.time_host_pages(op, pages, args.steps, small)
        r["page_rows"] = pr
        r["gb_per_s"] = r["rows_per_s"] * host.bytes_per_row / 1e9
        out.setdefault("add_page", []).append(r)
        print(json.dumps(r), file=sys.stderr)
    if args.serialized:
        for pr in (1_000_000, 65536):
            rows = min(args.rows, 32_000_000)
            payloads = host.serialized_pages(pr, rows)
            r = B.time_serialized_pages(op, payloads, args.steps, small)
            r["page_rows"] = pr
            r["gb_per_s"] = r["rows_per_s"] * host.bytes_per_row / 1e9
            out.setdefault("add_serialized_page", []).append(r)
            print(json.dumps(r), file=sys.stderr)
    op.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
