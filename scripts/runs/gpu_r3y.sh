#!/bin/bash
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5) | tee gpurun_out/r3y_pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r3y_bench_1gpu.json 2> gpurun_out/r3y_bench_1gpu.err || tail -20 gpurun_out/r3y_bench_1gpu.err
python scripts/bench_brief.py "N=1" < gpurun_out/r3y_bench_1gpu.json
