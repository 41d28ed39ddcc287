#!/bin/bash
# usage: gpu_rNgpu.sh N tag
N=$1; TAG=$2
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus $N > gpurun_out/${TAG}_bench_${N}gpu.json 2> gpurun_out/${TAG}_bench_${N}gpu.err || tail -20 gpurun_out/${TAG}_bench_${N}gpu.err
python scripts/bench_brief.py "N=$N" < gpurun_out/${TAG}_bench_${N}gpu.json
