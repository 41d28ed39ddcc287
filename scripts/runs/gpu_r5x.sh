#!/bin/bash
# round 2, call x: ncu --set full of the K3 GEMM kernel: fold block (32 B swizzle) + early TMEM release
mkdir -p gpurun_out
CMD2="python scripts/exact_probe.py 1000000,128,10000,10,0"
$CMD2 > gpurun_out/r5x_exact_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:exact_gemm_kernel --launch-skip 1 --launch-count 1 -f -o gpurun_out/r5x_exact_gemm_full $CMD2 > gpurun_out/r5x_ncu_exact.log 2>&1
tail -3 gpurun_out/r5x_exact_plain.log; tail -2 gpurun_out/r5x_ncu_exact.log
ls -la gpurun_out/ | grep r5n
