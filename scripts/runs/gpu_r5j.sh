#!/bin/bash
# round 2, call j: tests on the NANG split + direct-in queries; default bench; A/B of the page-locked-query path
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) | tee gpurun_out/r5j_pytest_gpu.log
SECONDS=0
timeout 1500 python bench.py --steps 20 --warmup 5 > gpurun_out/r5j_bench_1gpu.json 2> gpurun_out/r5j_bench_1gpu.err; echo "full rc=$? wall=${SECONDS}s"; tail -5 gpurun_out/r5j_bench_1gpu.err
python scripts/bench_brief.py "N=1" < gpurun_out/r5j_bench_1gpu.json
for v in 1 0 1 0; do
  B200HNSW_DIRECT_IN=$v python bench.py --steps 50 --warmup 5 --sides none --no-cpu-baseline 2>/dev/null | python scripts/bench_brief.py "direct_in=$v" | head -1
done | tee gpurun_out/r5j_direct_in_ab.log
