#!/bin/bash
# round 2, call 8p: full GPU suite on the new Add schedule + the default bench run with build.reference_order_at_full_size
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/r8p_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/r8p_pytest_gpu.log | cut -c1-300
timeout 1500 python bench.py --steps 20 --warmup 5 > gpurun_out/r8p_bench_1gpu.json 2> gpurun_out/r8p_bench_1gpu.err
echo "bench rc=$?"; timeout 20 python scripts/bench_brief.py r8p < gpurun_out/r8p_bench_1gpu.json | head -30
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r8p_bench_1gpu.json").read().strip().splitlines()[-1])
print(json.dumps(d["build"], indent=1)[:3000])
PY
