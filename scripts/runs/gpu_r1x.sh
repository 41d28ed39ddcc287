#!/bin/bash
mkdir -p gpurun_out
CMD="python scripts/k1_variants.py 1000000 0,1,9,11,G16 10000,100000"
$CMD > gpurun_out/k1_variants_r1x.log 2>&1; echo rc=$?; cat gpurun_out/k1_variants_r1x.log
CMD="python scripts/k1_variants.py 1000000 0 10000"
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:traverse_kernel -s 3 -c 1 -f -o gpurun_out/r1x_k1v4_full $CMD > gpurun_out/ncu_full_r1x.log 2>&1
tail -3 gpurun_out/ncu_full_r1x.log
