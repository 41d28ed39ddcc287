#!/bin/bash
# round 2, call 8g: K2s with workers keeping their speculation current while it waits (claim handshake): parity, timings
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_add_window_gpu.py tests/test_add_parity_gpu.py -q -m gpu > gpurun_out/r8g_pytest.log 2>&1
echo "tests rc=$?"; tail -8 gpurun_out/r8g_pytest.log | cut -c1-300
timeout 900 python scripts/k2s_probe.py 1000000 128 8000 8,16,32,48,64,96 > gpurun_out/r8g_k2s_1m.log 2>&1
cat gpurun_out/r8g_k2s_1m.log
timeout 900 python scripts/k2s_probe.py 100000 128 4000 4,8,16,32 > gpurun_out/r8g_k2s_100k.log 2>&1
cat gpurun_out/r8g_k2s_100k.log
