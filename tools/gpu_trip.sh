#!/bin/bash
mkdir -p gpurun_out
exec > gpurun_out/trip.log 2>&1
timeout 900 python -m pytest tests/test_gpu_exchange.py -m gpu -x -q 2>&1 | tail -15
