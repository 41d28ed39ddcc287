#!/bin/bash
# round 2, call 7f: fresh ncu --set full captures of the kernels as shipped at the end of the round:
# K1 traverse_kernel<8,4,2,10,2,NANG=0> (1 M x 128 L2, 10 k queries) and K1x traverse_ext_kernel (ef = 64)
mkdir -p gpurun_out
CMD="python scripts/k1_variants.py 1000000 0 10000"
$CMD > gpurun_out/r7f_k1_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:traverse_kernel -s 3 -c 1 -f -o gpurun_out/r7f_k1_full $CMD > gpurun_out/r7f_ncu_k1.log 2>&1
tail -1 gpurun_out/r7f_k1_plain.log; tail -1 gpurun_out/r7f_ncu_k1.log
CMD2="python scripts/k1x_probe.py 1000000 uniform 64 10000"
$CMD2 > gpurun_out/r7f_k1x_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:traverse_ext_kernel -s 2 -c 1 -f -o gpurun_out/r7f_k1x_full $CMD2 > gpurun_out/r7f_ncu_k1x.log 2>&1
tail -2 gpurun_out/r7f_k1x_plain.log; tail -1 gpurun_out/r7f_ncu_k1x.log
