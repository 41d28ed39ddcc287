#!/bin/bash
# round 2, call 6c (2 GPUs): negative_test.go replay; reduced N = 2 bench after the callers / cpu_baseline change
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_negatives.py -x -q 2>&1 | tail -6) | tee gpurun_out/r6c_pytest_negatives.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
SECONDS=0
timeout 900 $TR bench.py --gpus 2 --rows 200000 --steps 10 --cpu-seconds 2 > gpurun_out/r6c_bench_2gpu_small.json 2> gpurun_out/r6c_bench_2gpu_small.err; echo "small rc=$? wall=${SECONDS}s"; tail -5 gpurun_out/r6c_bench_2gpu_small.err
python scripts/bench_brief.py "N=2 small" < gpurun_out/r6c_bench_2gpu_small.json
