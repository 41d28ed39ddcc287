#!/bin/bash
mkdir -p gpurun_out
CMD="python scripts/exact_probe.py 1000000,128,10000,10,0"
$CMD > gpurun_out/exact_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:exact_gemm_kernel -s 1 -c 1 -f -o gpurun_out/r1p_gemm_full $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
