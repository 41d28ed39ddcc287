#!/bin/bash
# round 2, call l: the no-allocation ExactIndex entry point (alone in a fresh process, then behind every other test);
# default bench with the extended-build block of the clustered side
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q 2>&1 | tail -15) | tee gpurun_out/r5l_pytest_noalloc_alone.log
(timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) | tee gpurun_out/r5l_pytest_gpu.log
SECONDS=0
timeout 1800 python bench.py --steps 20 --warmup 5 > gpurun_out/r5l_bench_1gpu.json 2> gpurun_out/r5l_bench_1gpu.err; echo "full rc=$? wall=${SECONDS}s"; tail -5 gpurun_out/r5l_bench_1gpu.err
python scripts/bench_brief.py "N=1" < gpurun_out/r5l_bench_1gpu.json
