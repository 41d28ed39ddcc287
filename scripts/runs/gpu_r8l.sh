#!/bin/bash
# round 2, call 8l: insertion kernels with the block-wide candidate scan, warp-wide eviction decision, own list in shared memory
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_add_window_gpu.py tests/test_add_parity_gpu.py tests/test_delete_gpu.py tests/test_dropin_gpu.py tests/test_wave_add_gpu.py -q -m gpu > gpurun_out/r8l_pytest.log 2>&1
echo "tests rc=$?"; tail -8 gpurun_out/r8l_pytest.log | cut -c1-300
B200HNSW_TRACE_ADD=1 timeout 900 python scripts/k2s_probe.py 1000000 128 8000 1,32 > gpurun_out/r8l_k2s_1m.log 2>&1
cat gpurun_out/r8l_k2s_1m.log
