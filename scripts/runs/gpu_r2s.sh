#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_exact_gpu.py -m gpu -x -q 2>&1 | tail -2
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2s_exact_launches.csv python scripts/exact_probe.py 1000000,128,10000,10,0 250000,1024,4096,10,1 > gpurun_out/ncu_r2s.log 2>&1
tail -2 gpurun_out/ncu_r2s.log
