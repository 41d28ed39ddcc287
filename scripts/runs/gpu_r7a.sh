#!/bin/bash
# round 2, call 7a: final state on one GPU: smoke, full suite, default bench, reference arm
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3 | tee gpurun_out/r7a_smoke.log
(timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) | tee gpurun_out/r7a_pytest_gpu.log
SECONDS=0
timeout 1800 python bench.py --steps 20 --warmup 5 > gpurun_out/r7a_bench_1gpu.json 2> gpurun_out/r7a_bench_1gpu.err; echo "full rc=$? wall=${SECONDS}s"; tail -5 gpurun_out/r7a_bench_1gpu.err
python scripts/bench_brief.py "N=1" < gpurun_out/r7a_bench_1gpu.json
SECONDS=0
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r7a_bench_ref_1gpu.json 2> gpurun_out/r7a_bench_ref_1gpu.err; echo "ref rc=$? wall=${SECONDS}s"; head -c 500 gpurun_out/r7a_bench_ref_1gpu.json
