#!/bin/bash
# round 2, call h (2 GPUs): the multi-GPU tests (skipped on 1-GPU boxes), then bench.py at N = 2: reduced, then full default
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q 2>&1 | tail -5) | tee gpurun_out/r5h_pytest_2gpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
SECONDS=0
timeout 900 $TR bench.py --gpus 2 --rows 100000 --steps 10 --cpu-seconds 2 > gpurun_out/r5h_bench_2gpu_small.json 2> gpurun_out/r5h_bench_2gpu_small.err; echo "small rc=$? wall=${SECONDS}s"; tail -5 gpurun_out/r5h_bench_2gpu_small.err
python scripts/bench_brief.py "N=2 small" < gpurun_out/r5h_bench_2gpu_small.json
SECONDS=0
timeout 1500 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r5h_bench_2gpu.json 2> gpurun_out/r5h_bench_2gpu.err; echo "full rc=$? wall=${SECONDS}s"; tail -5 gpurun_out/r5h_bench_2gpu.err
python scripts/bench_brief.py "N=2" < gpurun_out/r5h_bench_2gpu.json
