#!/bin/bash
# round 2, call 8r: replenish continues on the list the back-link removal has just loaded (one round trip less per eviction)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_add_window_gpu.py tests/test_add_parity_gpu.py tests/test_delete_gpu.py -q -m gpu -x > gpurun_out/r8r_pytest.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/r8r_pytest.log | cut -c1-200
B200HNSW_TRACE_ADD=1 timeout 120 python scripts/k2s_probe.py 1000000 128 8000 1,32,32 > gpurun_out/r8r_k2s_1m.log 2>&1
cat gpurun_out/r8r_k2s_1m.log
