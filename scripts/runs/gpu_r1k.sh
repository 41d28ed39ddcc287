#!/bin/bash
mkdir -p gpurun_out
CMD="python scripts/k1_variants.py 1000000 0 10000"
$CMD > gpurun_out/k1v3_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:traverse_kernel -s 3 -c 1 -f -o gpurun_out/r1k_k1v3_full $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
