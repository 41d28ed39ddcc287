#!/bin/bash
# round 2, call 8f: block-wide layer search inside the insertion kernels (K2 and K2s): parity, then timings
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_add_window_gpu.py tests/test_add_parity_gpu.py tests/test_delete_gpu.py tests/test_dropin_gpu.py -q -m gpu > gpurun_out/r8f_pytest.log 2>&1
echo "tests rc=$?"; tail -8 gpurun_out/r8f_pytest.log | cut -c1-300
timeout 900 python scripts/k2s_probe.py 1000000 128 8000 1,8,16,24,0 > gpurun_out/r8f_k2s_1m.log 2>&1
cat gpurun_out/r8f_k2s_1m.log
