#!/bin/bash
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5) | tee gpurun_out/r3o_pytest_gpu.log
timeout 900 python bench.py --no-exact > gpurun_out/r3o_bench_1gpu.json 2> gpurun_out/r3o_bench_1gpu.err || tail -20 gpurun_out/r3o_bench_1gpu.err
python - <<'PY'
import json
l=json.load(open('gpurun_out/r3o_bench_1gpu.json'))
print({k:l[k] for k in ('value','ms_per_step','build','parity')})
PY
