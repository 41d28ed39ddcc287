#!/bin/bash
# round 2, call g: what does the NaN guard cost K1?  A/B of the shipped library against a build with the guard compiled out
mkdir -p gpurun_out
for rep in 1 2 3; do
  python scripts/k1_variants.py 1000000 0 10000 2>&1 | tail -1 | sed 's/^/guard   /'
  B200HNSW_LIB=hnsw_b200/lib/variants/libb200hnsw_noguard.so python scripts/k1_variants.py 1000000 0 10000 2>&1 | tail -1 | sed 's/^/noguard /'
done | tee gpurun_out/r5g_k1_guard_ab.log
for rep in 1 2; do
  python bench.py --steps 50 --warmup 5 --sides none --no-cpu-baseline 2>/dev/null | python scripts/bench_brief.py guard | head -1
  B200HNSW_LIB=hnsw_b200/lib/variants/libb200hnsw_noguard.so python bench.py --steps 50 --warmup 5 --sides none --no-cpu-baseline 2>/dev/null | python scripts/bench_brief.py noguard | head -1
done | tee -a gpurun_out/r5g_k1_guard_ab.log
