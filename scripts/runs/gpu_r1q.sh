#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_exact_gpu.py -m gpu -x -q > gpurun_out/pytest_exact.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_exact.log
timeout 900 python scripts/exact_probe.py 1000000,128,10000,10,0 250000,1024,4096,10,1 1000000,1024,64,10,1 2>&1 | tee gpurun_out/exact_probe3.log
