#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_exact_gpu.py -m gpu -x -q -k "tensor" > gpurun_out/pytest_exact_tensor.log 2>&1; echo "pytest rc=$?"; tail -40 gpurun_out/pytest_exact_tensor.log
nvidia-smi --query-gpu=name,memory.used --format=csv
