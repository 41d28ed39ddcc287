#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_exact_gpu.py -m gpu -x -q -k "not tensor" > gpurun_out/pytest_exact.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_exact.log
