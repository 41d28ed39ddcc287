#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
for c in 1 2 4 8; do B200HNSW_E2E_CHUNKS=$c timeout 600 python bench.py --no-cpu-baseline --steps 100 > gpurun_out/bench_c$c.json 2>gpurun_out/bench_c$c.err; python - <<PY
import json
d=json.load(open('gpurun_out/bench_c$c.json'))
print('chunks $c value', d['value'], 'e2e', d['e2e']['value'], 'kernel_ms', d['roofline']['kernel_ms'], 'frac', d['roofline']['frac'])
PY
done
timeout 600 python bench.py > gpurun_out/bench_1m.json 2>gpurun_out/bench_1m.err; tail -c 1800 gpurun_out/bench_1m.json
