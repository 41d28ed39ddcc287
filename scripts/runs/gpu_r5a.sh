#!/bin/bash
# round 2, call a: baseline on this pod + fresh ncu of the SHIPPED K1 template and tensor-pipe counters for K3
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv | tee gpurun_out/r5a_smi.txt
(timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5) | tee gpurun_out/r5a_pytest_gpu.log
CMD="python scripts/k1_variants.py 1000000 0 10000"
$CMD > gpurun_out/r5a_k1_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:traverse_kernel -s 3 -c 1 -f -o gpurun_out/r5a_k1_full $CMD > gpurun_out/r5a_ncu_k1.log 2>&1
tail -2 gpurun_out/r5a_k1_plain.log; tail -2 gpurun_out/r5a_ncu_k1.log
CMD2="python scripts/exact_probe.py 1000000,128,10000,10,0"
$CMD2 > gpurun_out/r5a_exact_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:exact_gemm_kernel --launch-skip 1 --launch-count 1 -f -o gpurun_out/r5a_exact_gemm_full $CMD2 > gpurun_out/r5a_ncu_exact.log 2>&1
tail -3 gpurun_out/r5a_exact_plain.log; tail -2 gpurun_out/r5a_ncu_exact.log
ncu --query-metrics 2>/dev/null | grep -i -E "tensor|tma|utc" | head -80 > gpurun_out/r5a_tensor_metric_names.txt
ls -la gpurun_out/ | grep r5a
