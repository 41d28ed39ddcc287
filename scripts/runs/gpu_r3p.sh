#!/bin/bash
# final state of the round: reference arm, default bench
mkdir -p gpurun_out
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r3p_bench_reference_arm.json 2> gpurun_out/r3p_ref.err || tail -5 gpurun_out/r3p_ref.err
cut -c1-260 gpurun_out/r3p_bench_reference_arm.json
timeout 900 python bench.py > gpurun_out/r3p_bench_1gpu.json 2> gpurun_out/r3p_bench_1gpu.err || tail -20 gpurun_out/r3p_bench_1gpu.err
python - <<'PY'
import json
l=json.load(open('gpurun_out/r3p_bench_1gpu.json'))
print({k:l[k] for k in ('value','ms_per_step','e2e','roofline','parity','clocks','build','recall_at_10','cpu_baseline')})
print(l['exact'])
PY
