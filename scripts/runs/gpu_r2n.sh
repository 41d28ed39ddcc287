#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_sharded_gpu.py tests/test_search_parity_gpu.py -m gpu -x -q 2>&1 | tail -3
for N in 8 4 2; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --steps 50 --warmup 3 > gpurun_out/bench_g${N}_r2n.json 2> gpurun_out/bench_g${N}_r2n.err; echo "N=$N rc=$?"; cut -c1-230 gpurun_out/bench_g${N}_r2n.json; grep -o '"e2e": {[^}]*}' gpurun_out/bench_g${N}_r2n.json; grep -o '"kernel_ms": [0-9.]*' gpurun_out/bench_g${N}_r2n.json; tail -1 gpurun_out/bench_g${N}_r2n.err
done
