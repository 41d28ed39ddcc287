#!/bin/bash
mkdir -p gpurun_out
(timeout 600 python scripts/e2e_callers_probe.py 2>&1 | grep -v Warning; CUDA_DEVICE_MAX_CONNECTIONS=32 timeout 600 python scripts/e2e_callers_probe.py 2>&1 | grep -v Warning) | tee gpurun_out/r3h_e2e_callers.log
