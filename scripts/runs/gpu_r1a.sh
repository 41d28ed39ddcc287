#!/bin/bash
# round-1 first GPU pass: parity tests, probes, small bench, ncu launch list + full capture of K1
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python scripts/k2_probe.py 20000 128 16 64 > gpurun_out/k2_probe.log 2>&1; tail -12 gpurun_out/k2_probe.log
timeout 600 python scripts/k1_probe.py 200000 128 16 64 10000,100000,1000000 > gpurun_out/k1_probe.log 2>&1; tail -20 gpurun_out/k1_probe.log
timeout 600 python bench.py --n 20000 --steps 5 --warmup 3 > gpurun_out/bench_20k.log 2>&1; tail -3 gpurun_out/bench_20k.log
