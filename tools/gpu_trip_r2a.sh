#!/bin/bash
# round 2, visit a: the coalescing ingest path — parity tests, then rows/s at the reference's page shapes against the H2D ceiling
mkdir -p gpurun_out
exec > gpurun_out/r2a.log 2>&1
set -x
date +%s
nproc; free -g | head -2; lscpu | grep -i "model name\|socket\|numa\|thread(s)\|L3"; nvidia-smi topo -m | head -20
timeout 600 python -m pytest tests/test_gpu_ingest.py -q --timeout 120 -p no:cacheprovider > gpurun_out/r2a_pytest_ingest.log 2>&1; echo "pytest ingest exit $?"
tail -30 gpurun_out/r2a_pytest_ingest.log
date +%s
for t in 16 8 4 1; do
  PGPU_COPY_THREADS=$t timeout 300 python tools/ingest_bench.py --rows 128000000 --page-rows 1000000,65536 --steps 3 > gpurun_out/r2a_ingest_t$t.json 2> gpurun_out/r2a_ingest_t$t.err; echo "ingest t=$t exit $?"
  cat gpurun_out/r2a_ingest_t$t.json; tail -3 gpurun_out/r2a_ingest_t$t.err
done
PGPU_COPY_THREADS=8 PGPU_STAGE_MB=16 timeout 300 python tools/ingest_bench.py --rows 128000000 --page-rows 1000000,65536,4096 --steps 3 --serialized > gpurun_out/r2a_ingest_s16.json 2> gpurun_out/r2a_ingest_s16.err; cat gpurun_out/r2a_ingest_s16.json; tail -3 gpurun_out/r2a_ingest_s16.err
PGPU_COPY_THREADS=8 PGPU_STAGE_MB=256 timeout 300 python tools/ingest_bench.py --rows 128000000 --page-rows 1000000,65536 --steps 3 > gpurun_out/r2a_ingest_s256.json 2> gpurun_out/r2a_ingest_s256.err; cat gpurun_out/r2a_ingest_s256.json; tail -3 gpurun_out/r2a_ingest_s256.err
date +%s
timeout 900 python -m pytest tests -m gpu -q -x --timeout 120 -p no:cacheprovider > gpurun_out/r2a_pytest_gpu.log 2>&1; echo "pytest gpu exit $?"
tail -15 gpurun_out/r2a_pytest_gpu.log
date +%s
