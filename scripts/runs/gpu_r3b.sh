#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_hybrid.py -m gpu -x -q > gpurun_out/r3b_hybrid.log 2>&1; tail -15 gpurun_out/r3b_hybrid.log
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r3b_all.log 2>&1; tail -5 gpurun_out/r3b_all.log
