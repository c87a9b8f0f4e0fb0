#!/bin/bash
# short visit: GPU tests of the serialized-page ingest, then the ingest comparison
mkdir -p gpurun_out
exec > gpurun_out/trip3.log 2>&1
set -x
timeout 200 python -m pytest tests/test_pages_serde.py -m gpu -x -q --timeout 120 2>&1 | tail -30
timeout 60 python tools/serde_ingest_bench.py
