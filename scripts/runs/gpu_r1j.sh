#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/e2e_probe.py 1000000 10000 1,2,4,8 2>&1 | tee gpurun_out/e2e_probe.log
