#!/bin/bash
# round 2, call e: tests again (drop-in + extended build), K1x probe + ncu --set full of the extended-mode kernel
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) | tee gpurun_out/r5e_pytest_gpu.log
python scripts/k1x_probe.py 1000000 uniform 16,64,256 10000 > gpurun_out/r5e_k1x_uniform.log 2>&1; tail -4 gpurun_out/r5e_k1x_uniform.log
python scripts/k1x_probe.py 1000000 clustered 16,64,256 10000 1 > gpurun_out/r5e_k1x_clustered_extbuild.log 2>&1; tail -4 gpurun_out/r5e_k1x_clustered_extbuild.log
CMD="python scripts/k1x_probe.py 1000000 uniform 64 10000"
$CMD > gpurun_out/r5e_k1x_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:traverse_ext_kernel -s 2 -c 1 -f -o gpurun_out/r5e_k1x_full $CMD > gpurun_out/r5e_ncu_k1x.log 2>&1
tail -2 gpurun_out/r5e_ncu_k1x.log
