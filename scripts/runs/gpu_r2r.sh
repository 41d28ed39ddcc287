#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_exact_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 900 python scripts/exact_probe.py 1000000,128,10000,10,0 250000,1024,4096,10,1 2>&1 | grep -v "it=0" | tee gpurun_out/exact_probe_r2r.log
