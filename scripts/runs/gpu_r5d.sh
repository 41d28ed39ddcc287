#!/bin/bash
# round 2, call d: drop-in tests (replace existing key, Export after Delete, string/float keys) + full default bench on 1 GPU
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) | tee gpurun_out/r5d_pytest_gpu.log
SECONDS=0
timeout 1500 python bench.py --steps 20 --warmup 5 > gpurun_out/r5d_bench_1gpu.json 2> gpurun_out/r5d_bench_1gpu.err; echo "full rc=$? wall=${SECONDS}s"; tail -5 gpurun_out/r5d_bench_1gpu.err
python scripts/bench_brief.py "N=1" < gpurun_out/r5d_bench_1gpu.json
