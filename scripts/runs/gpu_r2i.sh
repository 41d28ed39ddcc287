#!/bin/bash
mkdir -p gpurun_out
export B200HNSW_LIB=$PWD/hnsw_b200/lib/exp_T.so
timeout 600 python scripts/k1_bscale.py 1000000 > gpurun_out/k1_timing_r2i.log 2>&1; wc -l gpurun_out/k1_timing_r2i.log
