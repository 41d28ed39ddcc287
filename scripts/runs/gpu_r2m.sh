#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 600 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q 2>&1 | tail -3
for N in 8 4; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --steps 50 --warmup 3 > gpurun_out/bench_g${N}_r2m.json 2> gpurun_out/bench_g${N}_r2m.err; echo "N=$N rc=$?"; cut -c1-230 gpurun_out/bench_g${N}_r2m.json; grep -o '"e2e": {[^}]*}' gpurun_out/bench_g${N}_r2m.json; tail -2 gpurun_out/bench_g${N}_r2m.err
done
