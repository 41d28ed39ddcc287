#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/k1_variants.py 1000000 0,G16,G32,1,2,3,4,5,6,7 10000,100000 > gpurun_out/k1_variants.log 2>&1
cat gpurun_out/k1_variants.log
