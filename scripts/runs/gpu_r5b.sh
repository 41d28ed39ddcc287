#!/bin/bash
# round 2, call b: ADVICE fixes (NaN-safe queue, visited retry, load/delete consistency) + extended mode K1x parity
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) | tee gpurun_out/r5b_pytest_gpu.log
