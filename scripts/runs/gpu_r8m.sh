#!/bin/bash
# round 2, call 8m: cycles per phase of the sequential insertion kernel (exploration build, -DB200_K2_TIMING)
mkdir -p gpurun_out
timeout 600 python scripts/k2s_probe.py 1000000 128 2000 1 > gpurun_out/r8m_k2_phases_1m.log 2>&1; cat gpurun_out/r8m_k2_phases_1m.log
timeout 600 python scripts/k2s_probe.py 20000 128 2000 1 > gpurun_out/r8m_k2_phases_20k.log 2>&1; cat gpurun_out/r8m_k2_phases_20k.log
