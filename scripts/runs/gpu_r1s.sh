#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 900 python bench.py > gpurun_out/bench_r1s.json 2> gpurun_out/bench_r1s.err; echo "rc=$?"; tail -c 6000 gpurun_out/bench_r1s.json; tail -5 gpurun_out/bench_r1s.err
