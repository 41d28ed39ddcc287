#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python scripts/wave_probe.py 200000 128 16 64 4096 10000,100000 > gpurun_out/wave_probe_200k.log 2>&1; tail -12 gpurun_out/wave_probe_200k.log
for G in 8 16 32; do B200HNSW_K1_G=$G timeout 600 python scripts/wave_probe.py 100000 128 16 64 4096 10000,100000 > gpurun_out/wave_probe_G$G.log 2>&1; tail -4 gpurun_out/wave_probe_G$G.log; done
