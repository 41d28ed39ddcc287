#!/bin/bash
# round 2, call 6h: pageable queries copied by helper threads: robustness tests + e2e figures (3 runs)
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_robustness_gpu.py tests/test_search_parity_gpu.py -x -q 2>&1 | tail -5) | tee gpurun_out/r6h_pytest.log
for i in 1 2 3; do
  python bench.py --steps 50 --warmup 5 --sides none --no-cpu-baseline 2>/dev/null | python scripts/bench_brief.py "run$i" 2>/dev/null | head -1
done | tee gpurun_out/r6h_pageable_threads.log
