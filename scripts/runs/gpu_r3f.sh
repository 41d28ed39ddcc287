#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python bench.py > gpurun_out/r3f_bench_1gpu.json 2> gpurun_out/r3f_bench_1gpu.err || { tail -20 gpurun_out/r3f_bench_1gpu.err; exit 1; }
cat gpurun_out/r3f_bench_1gpu.json | cut -c1-2500
