#!/bin/bash
mkdir -p gpurun_out
CMD="python scripts/exact_probe.py 250000,1024,4096,10,1"
$CMD > gpurun_out/exact_plain_r2o.log 2>&1; echo rc=$?; tail -5 gpurun_out/exact_plain_r2o.log
ncu --set full --clock-control none --import-source on -k regex:exact_gemm_kernel -s 1 -c 1 -f -o gpurun_out/r2o_exact_gemm_full $CMD > gpurun_out/ncu_full_r2o.log 2>&1
tail -2 gpurun_out/ncu_full_r2o.log
