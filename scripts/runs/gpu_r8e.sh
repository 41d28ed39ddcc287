#!/bin/bash
# round 2, call 8e: K2s after the 64-layer mask fix: window tests, the other Add tests, timings
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_add_window_gpu.py -q -m gpu > gpurun_out/r8e_pytest_window.log 2>&1
echo "window tests rc=$?"; tail -12 gpurun_out/r8e_pytest_window.log | cut -c1-300
timeout 900 python -m pytest tests/test_add_parity_gpu.py tests/test_delete_gpu.py tests/test_cpp_host.py tests/test_dropin_gpu.py -q -m gpu > gpurun_out/r8e_pytest_add.log 2>&1
echo "add tests rc=$?"; tail -3 gpurun_out/r8e_pytest_add.log
