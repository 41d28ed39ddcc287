#!/bin/bash
mkdir -p gpurun_out
CMD="python scripts/exact_probe.py 1000000,128,10000,10,0"
ncu --set full --clock-control none --import-source on -k regex:exact_gemm_kernel -s 1 -c 1 -f -o gpurun_out/r2q_exact_gemm128_full $CMD > gpurun_out/ncu_full_r2q.log 2>&1
tail -2 gpurun_out/ncu_full_r2q.log
