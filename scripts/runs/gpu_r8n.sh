#!/bin/bash
# round 2, call 8n: insertion search without the candidate queue once the result queue is full (K1's argument), prefetch
# hints for the next neighbour of the linking loop: parity, then timings
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_add_window_gpu.py tests/test_add_parity_gpu.py tests/test_delete_gpu.py tests/test_dropin_gpu.py tests/test_robustness_gpu.py -q -m gpu > gpurun_out/r8n_pytest.log 2>&1
echo "tests rc=$?"; tail -6 gpurun_out/r8n_pytest.log | cut -c1-300
B200HNSW_TRACE_ADD=1 timeout 900 python scripts/k2s_probe.py 1000000 128 8000 1,24,32,48 > gpurun_out/r8n_k2s_1m.log 2>&1
cat gpurun_out/r8n_k2s_1m.log
