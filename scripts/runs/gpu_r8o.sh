#!/bin/bash
# round 2, call 8o: window size vs rows for the automatic choice (20 k / 100 k / 300 k rows)
mkdir -p gpurun_out
timeout 600 python scripts/k2s_probe.py 20000 128 4000 1,2,4,8,16 > gpurun_out/r8o_k2s_20k.log 2>&1; cat gpurun_out/r8o_k2s_20k.log
timeout 600 python scripts/k2s_probe.py 100000 128 4000 1,8,16,32 > gpurun_out/r8o_k2s_100k.log 2>&1; cat gpurun_out/r8o_k2s_100k.log
timeout 600 python scripts/k2s_probe.py 300000 128 6000 8,16,32 > gpurun_out/r8o_k2s_300k.log 2>&1; cat gpurun_out/r8o_k2s_300k.log
timeout 600 python scripts/k2s_probe.py 5000 128 2000 1,2,4 > gpurun_out/r8o_k2s_5k.log 2>&1; cat gpurun_out/r8o_k2s_5k.log
