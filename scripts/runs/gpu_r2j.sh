#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/k1_variants.py 1000000 0,9,G16 10000,100000 > gpurun_out/k1_variants_r2j.log 2>&1
cat gpurun_out/k1_variants_r2j.log
export B200HNSW_LIB=$PWD/hnsw_b200/lib/exp_T.so
timeout 600 python scripts/k1_bscale.py 1000000 > gpurun_out/k1_timing_r2j.log 2>&1; wc -l gpurun_out/k1_timing_r2j.log
