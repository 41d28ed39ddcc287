#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/k1_variants.py 1000000 0,7,8,9,10,11,12,G16 10000,100000 > gpurun_out/k1_variants2.log 2>&1
cat gpurun_out/k1_variants2.log
CMD="python scripts/k1_variants.py 1000000 0 10000"
$CMD > gpurun_out/k1v2_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:traverse_kernel -s 3 -c 1 -f -o gpurun_out/r1f_k1v2_full $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
