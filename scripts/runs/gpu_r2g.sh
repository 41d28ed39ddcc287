#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2g.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_r2g.log
timeout 900 python scripts/k1_variants.py 1000000 0,9,13 10000,100000 > gpurun_out/k1_variants_r2g.log 2>&1
cat gpurun_out/k1_variants_r2g.log
timeout 600 python scripts/wave_probe.py 1000000 128 16 64 4096 10000 2>&1 | head -2
