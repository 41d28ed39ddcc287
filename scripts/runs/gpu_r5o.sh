#!/bin/bash
# round 2, call m: K3 epilogue v3/v4 (packed FMAs, group-wise re-score from TMEM, x16 loads, tail trims): tests, probe
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_exact_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -8) | tee gpurun_out/r5o_pytest_exact.log
timeout 900 python scripts/exact_probe.py 1000000,128,10000,10,0 250000,1024,4096,10,1 1000000,128,10000,10,1 1000000,96,10000,10,0 2>&1 | tee gpurun_out/r5o_exact_probe.log
