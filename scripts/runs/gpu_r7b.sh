#!/bin/bash
# round 2, call s (N GPUs, N = $1): multi-GPU tests (2 ranks, incl. search_all), then the default bench at N
N=${1:-4}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8; nproc
nvidia-smi topo -m 2>&1 | head -24 | tee gpurun_out/r7b_topo_${N}.txt
(timeout 900 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q 2>&1 | tail -5) | tee gpurun_out/r7b_pytest_2gpu_on_${N}.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
SECONDS=0
timeout 1500 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r7b_bench_${N}gpu.json 2> gpurun_out/r7b_bench_${N}gpu.err; echo "full rc=$? wall=${SECONDS}s"; tail -5 gpurun_out/r7b_bench_${N}gpu.err
python scripts/bench_brief.py "N=$N" < gpurun_out/r7b_bench_${N}gpu.json
SECONDS=0
timeout 600 $TR bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/r7b_bench_ref_${N}gpu.json 2> gpurun_out/r7b_bench_ref_${N}gpu.err; echo "ref rc=$? wall=${SECONDS}s"; head -c 600 gpurun_out/r7b_bench_ref_${N}gpu.json
