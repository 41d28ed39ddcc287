#!/bin/bash
# round 2, call c: the restructured bench.py (workloads, sides, e2e variants) on 1 GPU: reduced, then full default
mkdir -p gpurun_out
timeout 600 python bench.py --n 100000 --steps 10 --cpu-seconds 2 > gpurun_out/r5c_bench_small.json 2> gpurun_out/r5c_bench_small.err; echo "small rc=$?"; tail -5 gpurun_out/r5c_bench_small.err
python scripts/bench_brief.py "small" < gpurun_out/r5c_bench_small.json
/usr/bin/time -v timeout 1500 python bench.py --steps 20 --warmup 5 > gpurun_out/r5c_bench_1gpu.json 2> gpurun_out/r5c_bench_1gpu.err; echo "full rc=$?"; tail -25 gpurun_out/r5c_bench_1gpu.err | grep -E "Elapsed|Maximum resident|Error|error|Traceback" 
python scripts/bench_brief.py "N=1" < gpurun_out/r5c_bench_1gpu.json
