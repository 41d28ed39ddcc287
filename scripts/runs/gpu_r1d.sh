#!/bin/bash
set -x
mkdir -p gpurun_out
nproc; free -g | head -2
( time timeout 900 python bench.py > gpurun_out/bench_1m.json 2> gpurun_out/bench_1m.err ) 2>&1 | tail -3; tail -c 3000 gpurun_out/bench_1m.json; tail -5 gpurun_out/bench_1m.err
( time timeout 1200 python bench.py --impl reference --steps 20 > gpurun_out/bench_ref_1m.json 2> gpurun_out/bench_ref_1m.err ) 2>&1 | tail -3; tail -c 1500 gpurun_out/bench_ref_1m.json; tail -5 gpurun_out/bench_ref_1m.err
