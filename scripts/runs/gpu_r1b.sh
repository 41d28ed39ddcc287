#!/bin/bash
# ncu evidence for K1 on a 200k x 128 graph (oracle-built on the box's CPU): launch list + one full capture
set -x
mkdir -p gpurun_out
CMD="python scripts/k1_probe.py 200000 128 16 64 100000"
$CMD > gpurun_out/k1_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/r1b_k1_launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
$CMD > gpurun_out/k1_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:search_kernel -s 1 -c 1 -f -o gpurun_out/r1b_k1_full $CMD > gpurun_out/ncu_full.log 2>&1
tail -5 gpurun_out/ncu_full.log
ls -la gpurun_out
