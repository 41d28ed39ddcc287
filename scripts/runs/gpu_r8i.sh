#!/bin/bash
# round 2, call 8i: ncu source-level capture of the sequential insertion kernel (500 insertions into a 1 M-row graph)
mkdir -p gpurun_out
CMD="python scripts/k2s_probe.py 1000000 128 500 1"
$CMD > gpurun_out/r8i_plain.log 2>&1 && tail -1 gpurun_out/r8i_plain.log &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:insert_kernel -c 1 -f -o gpurun_out/r8i_k2_full $CMD > gpurun_out/r8i_ncu.log 2>&1
tail -2 gpurun_out/r8i_ncu.log; ls -la gpurun_out/r8i_k2_full.ncu-rep
