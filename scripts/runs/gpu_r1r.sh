#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 2 --steps 50 --warmup 3 > gpurun_out/bench_g2.json 2> gpurun_out/bench_g2.err; echo "rc=$?"; tail -c 1500 gpurun_out/bench_g2.json; tail -5 gpurun_out/bench_g2.err
timeout 600 python -m pytest tests -m gpu -x -q -k "merge_kernel" 2>&1 | tail -3
