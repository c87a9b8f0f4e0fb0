#!/bin/bash
mkdir -p gpurun_out
exec > gpurun_out/ncu_variants.log 2>&1
for v in "4" "0"; do
  PGPU_TUNE_RG=$v timeout 900 ncu --set full --clock-control none --import-source on -k regex:pgpu_k_page_staged -s 2 -c 1 -f -o gpurun_out/q1s_rg$v \
    python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --rows 200000000 2>&1 | tail -2
done
