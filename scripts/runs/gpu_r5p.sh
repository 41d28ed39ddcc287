#!/bin/bash
# round 2, calls p/q: K3 epilogue variants v5 (re-score from registers per half) and v6 (early TMEM release) against v4,
# with parts of the epilogue switched off (scripts/k3_dbg_probe.py; the B200EXACT_DBG switch existed only in those builds)
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_exact_gpu.py -x -q 2>&1 | tail -4) | tee gpurun_out/r5p_pytest_exact.log
timeout 600 python scripts/k3_dbg_probe.py 2>&1 | tee gpurun_out/r5p_k3_dbg.log
