#!/bin/bash
# round 2, call z: K3 as shipped (fold block, 32 B-swizzle tile in its own ring, early TMEM release): full suite, bench
mkdir -p gpurun_out
(timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -25) | tee gpurun_out/r5z_pytest_gpu.log
SECONDS=0
timeout 1800 python bench.py --steps 20 --warmup 5 > gpurun_out/r5z_bench_1gpu.json 2> gpurun_out/r5z_bench_1gpu.err; echo "full rc=$? wall=${SECONDS}s"; tail -5 gpurun_out/r5z_bench_1gpu.err
python scripts/bench_brief.py "N=1" < gpurun_out/r5z_bench_1gpu.json
