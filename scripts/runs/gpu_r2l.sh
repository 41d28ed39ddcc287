#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 2 --steps 50 --warmup 3 > gpurun_out/bench_g2_r2l.json 2> gpurun_out/bench_g2_r2l.err; echo "rc=$?"; cut -c1-330 gpurun_out/bench_g2_r2l.json; grep -o '"e2e": {[^}]*}' gpurun_out/bench_g2_r2l.json; grep -o '"kernel_ms": [0-9.]*' gpurun_out/bench_g2_r2l.json; tail -3 gpurun_out/bench_g2_r2l.err
