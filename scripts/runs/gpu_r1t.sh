#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r1t.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_r1t.log
timeout 900 python scripts/k1_variants.py 1000000 0,1,2,4,9,11,G16 10000,100000 > gpurun_out/k1_variants_r1t.log 2>&1
cat gpurun_out/k1_variants_r1t.log
