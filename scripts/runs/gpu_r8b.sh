#!/bin/bash
# round 2, call 8b: K2s from an empty graph with the trace on, then the window tests again
mkdir -p gpurun_out
timeout 300 python scripts/k2s_debug.py > gpurun_out/r8b_debug.log 2>&1; cat gpurun_out/r8b_debug.log | cut -c1-600
timeout 900 python -m pytest tests/test_add_window_gpu.py -q -m gpu > gpurun_out/r8b_pytest_window.log 2>&1
echo "window tests rc=$?"; tail -15 gpurun_out/r8b_pytest_window.log | cut -c1-300
timeout 600 python -m pytest tests/test_cpp_host.py -x -q -m gpu 2>&1 | tail -3
