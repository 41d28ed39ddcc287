#!/bin/bash
# round 2, call 6a: which queries of the new fold-path test differ
mkdir -p gpurun_out
timeout 600 python scripts/k3_fold_debug.py 2>&1 | tee gpurun_out/r6a_fold_debug.log
