#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r3j_bench_1gpu.json 2> gpurun_out/r3j_bench_1gpu.err || tail -20 gpurun_out/r3j_bench_1gpu.err
python - <<'PY'
import json
l=json.load(open('gpurun_out/r3j_bench_1gpu.json'))
print({k:l[k] for k in ('value','ms_per_step','e2e','roofline','parity','clocks')})
PY
timeout 900 python bench.py --inflight 1 --no-exact --no-cpu-baseline > gpurun_out/r3j_bench_1gpu_inflight1.json 2>/dev/null
cut -c1-200 gpurun_out/r3j_bench_1gpu_inflight1.json
