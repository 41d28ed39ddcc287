#!/bin/bash
mkdir -p gpurun_out
CMD="python scripts/exact_probe.py 1000000,128,10000,10,0"
$CMD > gpurun_out/exact_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r1o_exact_launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
cut -d, -f5,15 gpurun_out/r1o_exact_launches.csv | tail -40
timeout 600 python scripts/exact_probe.py 250000,1024,4096,10,1 1000000,1024,64,10,1 2>&1 | tee gpurun_out/exact_probe2.log
