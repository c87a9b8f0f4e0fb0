#!/bin/bash
# visit 5: wire-format output + device generator + golden-answer tests
mkdir -p gpurun_out
exec > gpurun_out/trip5.log 2>&1
set -x
timeout 250 python -m pytest tests/test_pages_serde.py tests/test_tpch_golden.py -m gpu -q --timeout 120 2>&1 | tail -40
