#!/bin/bash
# round 2, call 7c: K3 fold epilogue handles both blocks of a tile together (one vote, both slot reads in flight)
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_exact_gpu.py tests/test_zz_no_alloc.py tests/test_hybrid.py -x -q 2>&1 | tail -6) | tee gpurun_out/r7c_pytest_exact.log
timeout 600 python scripts/exact_probe.py 1000000,128,10000,10,0 1000000,96,10000,10,0 250000,1024,4096,10,0 2>&1 | tee gpurun_out/r7c_exact_probe.log
