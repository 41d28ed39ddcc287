#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2a.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu_r2a.log
CMD="python scripts/wave_probe.py 1000000 128 16 64 4096 10000"
$CMD > gpurun_out/wave_probe_r2a.log 2>&1; echo rc=$?; cat gpurun_out/wave_probe_r2a.log
