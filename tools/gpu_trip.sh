#!/bin/bash
mkdir -p gpurun_out
exec > gpurun_out/trip.log 2>&1
timeout 600 python -m pytest tests/test_gpu_exchange.py tests/test_abi.py -m gpu -x -q -p no:cacheprovider --timeout 180 > gpurun_out/trip_pytest.log 2>&1; tail -25 gpurun_out/trip_pytest.log
