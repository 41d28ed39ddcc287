#!/bin/bash
# round 2, call r: K3 v4 as shipped: full GPU suite, probe with 16 and 8 epilogue warps, default bench
mkdir -p gpurun_out
(timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) | tee gpurun_out/r5r_pytest_gpu.log
timeout 600 python scripts/exact_probe.py 1000000,128,10000,10,0 250000,1024,4096,10,1 2>&1 | tee gpurun_out/r5r_exact_probe.log
B200EXACT_EPI_WARPS=8 timeout 600 python scripts/exact_probe.py 1000000,128,10000,10,0 2>&1 | tee gpurun_out/r5r_exact_probe_ew8.log
SECONDS=0
timeout 1800 python bench.py --steps 20 --warmup 5 > gpurun_out/r5r_bench_1gpu.json 2> gpurun_out/r5r_bench_1gpu.err; echo "full rc=$? wall=${SECONDS}s"; tail -5 gpurun_out/r5r_bench_1gpu.err
python scripts/bench_brief.py "N=1" < gpurun_out/r5r_bench_1gpu.json
