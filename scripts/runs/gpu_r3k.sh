#!/bin/bash
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q 2>&1 | tail -8) | tee gpurun_out/r3k_pytest_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 > gpurun_out/r3k_bench_2gpu.json 2> gpurun_out/r3k_bench_2gpu.err || tail -20 gpurun_out/r3k_bench_2gpu.err
cut -c1-300 gpurun_out/r3k_bench_2gpu.json
python - <<'PY'
import json
l=json.load(open('gpurun_out/r3k_bench_2gpu.json'))
print({k:l[k] for k in ('value','ms_per_step','e2e','roofline')})
PY
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --inflight 1 > gpurun_out/r3k_bench_2gpu_inflight1.json 2>/dev/null
python - <<'PY'
import json
l=json.load(open('gpurun_out/r3k_bench_2gpu_inflight1.json'))
print('inflight 1:', {k:l[k] for k in ('value','ms_per_step','e2e')})
PY
