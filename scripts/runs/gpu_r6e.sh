#!/bin/bash
# round 2, call 6e: own staging for pageable host buffers: robustness + parity tests, e2e pageable / pinned figures
mkdir -p gpurun_out
(timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) | tee gpurun_out/r6e_pytest_gpu.log
for v in 0 1; do
  if [ $v = 1 ]; then export B200HNSW_NO_HOST_STAGING=1; else unset B200HNSW_NO_HOST_STAGING; fi
  python bench.py --steps 50 --warmup 5 --sides none --no-cpu-baseline 2>/dev/null | python scripts/bench_brief.py "driver_staging=$v" 2>/dev/null | head -1
done | tee gpurun_out/r6e_pageable.log
