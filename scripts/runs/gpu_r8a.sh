#!/bin/bash
# round 2, call 8a: the speculative windowed Add (K2s) -- parity tests first, then the existing Add / delete / C++ host
# tests on the new default schedule, then timings per window at 100 k and 1 M rows
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_add_window_gpu.py -x -q -m gpu > gpurun_out/r8a_pytest_window.log 2>&1
echo "window tests rc=$?"; tail -5 gpurun_out/r8a_pytest_window.log
timeout 900 python -m pytest tests/test_add_parity_gpu.py tests/test_delete_gpu.py tests/test_cpp_host.py tests/test_dropin_gpu.py -x -q -m gpu > gpurun_out/r8a_pytest_add.log 2>&1
echo "add tests rc=$?"; tail -3 gpurun_out/r8a_pytest_add.log
timeout 600 python scripts/k2s_probe.py 100000 128 4000 1,2,4,8,0 > gpurun_out/r8a_k2s_100k.log 2>&1
cat gpurun_out/r8a_k2s_100k.log
timeout 900 python scripts/k2s_probe.py 1000000 128 8000 1,4,8,16,32,0 > gpurun_out/r8a_k2s_1m.log 2>&1
cat gpurun_out/r8a_k2s_1m.log
