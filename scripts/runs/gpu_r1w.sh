#!/bin/bash
mkdir -p gpurun_out
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw,temperature.gpu,clocks_throttle_reasons.active --format=csv
timeout 900 python scripts/k1_variants.py 1000000 0,9 10000,100000 > gpurun_out/k1_variants_r1w.log 2>&1
cat gpurun_out/k1_variants_r1w.log
timeout 900 python bench.py --no-exact --no-cpu-baseline > gpurun_out/bench_r1w.json 2> gpurun_out/bench_r1w.err; echo "rc=$?"; cut -c1-1800 gpurun_out/bench_r1w.json
