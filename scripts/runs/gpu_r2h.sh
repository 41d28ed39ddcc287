#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/k1_bscale.py 1000000 > gpurun_out/k1_bscale_r2h.log 2>&1; cat gpurun_out/k1_bscale_r2h.log
timeout 600 python scripts/e2e_probe.py 1000000 10000 1,2,4,8 > gpurun_out/e2e_probe_r2h.log 2>&1; cat gpurun_out/e2e_probe_r2h.log
