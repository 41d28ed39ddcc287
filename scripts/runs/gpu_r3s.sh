#!/bin/bash
mkdir -p gpurun_out
for i in 1 2 3 4 5 6 7 8 9 10; do
  B200HNSW_TRACE_SLOW=1 timeout 600 python bench.py --no-exact --no-cpu-baseline 2> gpurun_out/r3s_err_$i.log | python scripts/bench_brief.py "run $i"
  grep "slow call" gpurun_out/r3s_err_$i.log | tail -n +5 | head -8
done | tee gpurun_out/r3s_e2e_trace.log
