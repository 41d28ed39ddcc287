#!/bin/bash
# round 2, call 6b: fold path with the out-of-window queries kept finite: exact tests, then the full suite
mkdir -p gpurun_out
timeout 600 python scripts/k3_fold_debug.py 2>&1 | tail -12 | tee gpurun_out/r6b_fold_debug.log
(timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -12) | tee gpurun_out/r6b_pytest_gpu.log
