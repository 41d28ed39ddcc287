#!/bin/bash
# round 2, call 8s (last of the round): the whole GPU suite and the default bench line on the final tree; every step bounded
mkdir -p gpurun_out
timeout 150 python -m pytest tests -q -m gpu > gpurun_out/r8s_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/r8s_pytest_gpu.log | cut -c1-200
timeout 170 python bench.py --steps 20 --warmup 5 > gpurun_out/r8s_bench_1gpu.json 2> gpurun_out/r8s_bench_1gpu.err
echo "bench rc=$?"; timeout 20 python scripts/bench_brief.py r8s < gpurun_out/r8s_bench_1gpu.json | head -4 | cut -c1-400
