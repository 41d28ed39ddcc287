#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_delete_gpu.py -m gpu -x -q > gpurun_out/r3a_delete.log 2>&1
timeout 900 python -m pytest tests/test_wave_add_gpu.py -m gpu -x -q > gpurun_out/r3a_wave.log 2>&1
grep -n "Error\|error\|Fatal\|FAILED\|passed\|failed" gpurun_out/r3a_delete.log | head -20
grep -n "B200Error\|FAILED\|passed\|failed" gpurun_out/r3a_wave.log | head
