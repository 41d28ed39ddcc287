#!/bin/bash
mkdir -p gpurun_out
timeout 1700 compute-sanitizer --tool memcheck --error-exitcode 9 --launch-timeout 120 python -m pytest tests/test_wave_add_gpu.py tests/test_exact_gpu.py tests/test_search_parity_gpu.py tests/test_negatives.py -m gpu -x -q -k "not 1m and not large" > gpurun_out/sanitizer_r2x.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/sanitizer_r2x.log
