#!/bin/bash
# round 2, call t (2 GPUs): the reworked 2-rank search_all leg; K3 with the per-row term folded into the GEMM (Euclidean)
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q 2>&1 | tail -40) | tee gpurun_out/r5t_pytest_2gpu.log
(timeout 900 python -m pytest tests/test_exact_gpu.py tests/test_zz_no_alloc.py tests/test_hybrid.py -x -q 2>&1 | tail -30) | tee gpurun_out/r5t_pytest_exact.log
timeout 600 python scripts/exact_probe.py 1000000,128,10000,10,0 1000000,96,10000,10,0 250000,1024,4096,10,0 2>&1 | tee gpurun_out/r5t_exact_probe.log
