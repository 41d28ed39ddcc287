#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/exact_probe.py 1000000,128,10000,10,0 > gpurun_out/r3e_probe.log 2>&1 || exit 1
tail -5 gpurun_out/r3e_probe.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:exact_gemm_kernel --launch-skip 1 --launch-count 1 \
  -o gpurun_out/r3e_exact_gemm -f python scripts/exact_probe.py 1000000,128,10000,10,0 > gpurun_out/r3e_ncu.log 2>&1
tail -3 gpurun_out/r3e_ncu.log
