#!/bin/bash
# round 2, call 6j: K3 with four 128-column accumulator half-stages (N = 128 MMAs): tests, probe
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_exact_gpu.py tests/test_zz_no_alloc.py tests/test_hybrid.py -x -q 2>&1 | tail -12) | tee gpurun_out/r6j_pytest_exact.log
timeout 600 python scripts/exact_probe.py 1000000,128,10000,10,0 250000,1024,4096,10,1 1000000,128,10000,10,1 2>&1 | tee gpurun_out/r6j_exact_probe.log
B200EXACT_NO_SEED=1 timeout 600 python scripts/exact_probe.py 1000000,128,10000,10,0 2>&1 | tee gpurun_out/r6j_exact_probe_noseed.log
