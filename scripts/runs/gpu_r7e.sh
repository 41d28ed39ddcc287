#!/bin/bash
# round 2, call 7e: cosine reductions (dot, norm) advanced together: full suite (bit-exactness), C3 probe before/after is r7d
mkdir -p gpurun_out
(timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) | tee gpurun_out/r7e_pytest_gpu.log
timeout 900 python scripts/config_probe.py 1000000 768 32 64 cosine 10000 8 2000 2>&1 | tail -1 | cut -c1-900 | tee gpurun_out/r7e_c3_probe.log
