#!/bin/bash
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q 2>&1 | tail -15) | tee gpurun_out/r3w_pytest_2gpu.log
for ex in p2p nccl; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 2 --exchange $ex > gpurun_out/r3w_bench_2gpu_$ex.json 2> gpurun_out/r3w_bench_2gpu_$ex.err || tail -20 gpurun_out/r3w_bench_2gpu_$ex.err
  python scripts/bench_brief.py "N=2 $ex" < gpurun_out/r3w_bench_2gpu_$ex.json
done
