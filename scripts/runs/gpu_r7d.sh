#!/bin/bash
# round 2, call 7d: ncu --set full of K1 on BASELINE configs[2] (1 M x 768 cosine, M = 32): the DRAM counters of the
# traversal kernel on a workload whose rows are 3 KB (the C2 capture shows 6 % of the HBM peak: hub rows from L2)
mkdir -p gpurun_out
CMD="python scripts/config_probe.py 1000000 768 32 64 cosine 10000 8 200"
$CMD > gpurun_out/r7d_c3_plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:traverse_kernel --profile-from-start off --launch-skip 2 --launch-count 1 -f -o gpurun_out/r7d_k1_c3_full $CMD > gpurun_out/r7d_ncu_c3.log 2>&1
tail -2 gpurun_out/r7d_c3_plain.log | cut -c1-400; tail -2 gpurun_out/r7d_ncu_c3.log
