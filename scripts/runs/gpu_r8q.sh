#!/bin/bash
# round 2, call 8q: K2s with one inlined body of the insertion per kernel (worker) + an out-of-line one (committer): parity, timing
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_add_window_gpu.py -q -m gpu -x > gpurun_out/r8q_pytest.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/r8q_pytest.log | cut -c1-200
B200HNSW_TRACE_ADD=1 timeout 120 python scripts/k2s_probe.py 1000000 128 8000 32,32 > gpurun_out/r8q_k2s_1m.log 2>&1
cat gpurun_out/r8q_k2s_1m.log
