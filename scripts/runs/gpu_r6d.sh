#!/bin/bash
# round 2, call 6d: poll+yield wait in the host-pointer call: e2e with 1-4 callers; new delete / negatives tests; suite
mkdir -p gpurun_out
(timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) | tee gpurun_out/r6d_pytest_gpu.log
for c in 1 2 3 4; do
  python bench.py --steps 50 --warmup 5 --sides none --no-cpu-baseline --e2e-callers $c 2>/dev/null | python scripts/bench_brief.py "callers=$c" | head -1
done | tee gpurun_out/r6d_callers.log
