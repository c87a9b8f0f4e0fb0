#!/bin/bash
# round 2, visit c: memcheck of the bucketed page kernel
mkdir -p gpurun_out
exec > gpurun_out/r2c.log 2>&1
set -x
timeout 600 compute-sanitizer --tool memcheck --print-limit 8 python -m pytest "tests/test_gpu_highcard_variants.py::test_variant_equals_oracle[clustered-1-B]" -q --timeout 500 -p no:cacheprovider > gpurun_out/r2c_memcheck.log 2>&1; echo "memcheck exit $?"
grep -v "^=========     at\|Host Frame\|^=========         in " gpurun_out/r2c_memcheck.log | head -60
timeout 600 python -m pytest tests/test_gpu_highcard_variants.py -m gpu -q --timeout 200 -p no:cacheprovider > gpurun_out/r2c_pytest.log 2>&1; echo "pytest exit $?"
tail -12 gpurun_out/r2c_pytest.log
