#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 900 python scripts/k1_variants.py 1000000 0,1,4,9,11,G16,G32 10000,100000 > gpurun_out/k1_variants3.log 2>&1
cat gpurun_out/k1_variants3.log
