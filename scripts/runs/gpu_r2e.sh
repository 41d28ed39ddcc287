#!/bin/bash
mkdir -p gpurun_out
CMD="python scripts/k1_variants.py 1000000 0 10000"
$CMD > gpurun_out/k1_plain_r2e.log 2>&1; echo rc=$?; tail -1 gpurun_out/k1_plain_r2e.log
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:traverse_kernel -s 3 -c 1 -f -o gpurun_out/r2e_k1v4_full $CMD > gpurun_out/ncu_full_r2e.log 2>&1
tail -2 gpurun_out/ncu_full_r2e.log
