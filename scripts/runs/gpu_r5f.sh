#!/bin/bash
# round 2, call f: tests (drop-in, extended build, K1x v2 with W in registers) + K1x probe
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) | tee gpurun_out/r5f_pytest_gpu.log
python scripts/k1x_probe.py 1000000 uniform 16,64,256,512 10000 > gpurun_out/r5f_k1x_uniform.log 2>&1; tail -5 gpurun_out/r5f_k1x_uniform.log
python scripts/k1x_probe.py 1000000 clustered 16,64,256 10000 1 > gpurun_out/r5f_k1x_clustered_extbuild.log 2>&1; tail -4 gpurun_out/r5f_k1x_clustered_extbuild.log
python scripts/k1_variants.py 1000000 0 10000 2>&1 | tail -1
