#!/bin/bash
mkdir -p gpurun_out
for d in 2 3 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2957$d bench.py --gpus 2 --exchange p2p --inflight $d > gpurun_out/r3x_bench_2gpu_p2p_d$d.json 2> gpurun_out/r3x_err_$d.log || tail -20 gpurun_out/r3x_err_$d.log
  python scripts/bench_brief.py "N=2 p2p depth $d" < gpurun_out/r3x_bench_2gpu_p2p_d$d.json
done
