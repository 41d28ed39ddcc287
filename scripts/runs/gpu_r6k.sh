#!/bin/bash
# round 2, call 6k: per-phase cycle counts inside the K3 GEMM kernel (build with -DB200_K3_TIMING)
mkdir -p gpurun_out
timeout 600 python scripts/exact_probe.py 1000000,128,10000,10,0 2>&1 | grep -v "^$" | tail -40 | tee gpurun_out/r6k_k3_phase_cycles.log
