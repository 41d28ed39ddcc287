#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/k1_variants.py 1000000 0,13,14,15,16,11,G16 10000,100000 > gpurun_out/k1_variants_r2f.log 2>&1
cat gpurun_out/k1_variants_r2f.log
