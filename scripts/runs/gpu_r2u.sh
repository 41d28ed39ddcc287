#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 900 python bench.py > gpurun_out/bench_r2u.json 2> gpurun_out/bench_r2u.err; echo "rc=$?"; tail -3 gpurun_out/bench_r2u.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref_r2u.json 2> gpurun_out/bench_ref_r2u.err; echo "rc=$?"; cut -c1-600 gpurun_out/bench_ref_r2u.json
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
