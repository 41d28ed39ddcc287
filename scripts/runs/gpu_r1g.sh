#!/bin/bash
mkdir -p gpurun_out
CMD="python scripts/k1_variants.py 1000000 0 10000"
$CMD > gpurun_out/k1v2_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:traverse_kernel -s 3 -c 1 -f -o gpurun_out/r1g_k1v2_full $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 12 --csv --log-file gpurun_out/r1g_k1v2_launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
