#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/e2e_probe.py 1000000 10000 1,2,4 2>&1 | tail -5
B200HNSW_NO_DIRECT_OUT=1 timeout 600 python scripts/e2e_probe.py 1000000 10000 1,2,4 2>&1 | tail -5
timeout 900 python -m pytest tests/test_search_parity_gpu.py -m gpu -x -q 2>&1 | tail -2
