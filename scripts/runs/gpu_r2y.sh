#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29556 scripts/exact_sharded_probe.py 1000000,1024,10000,10,1 2>&1 | grep "exact sharded" | tee gpurun_out/exact_sharded_r2y.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29557 scripts/exact_sharded_probe.py 1000000,1024,64,10,1 2>&1 | grep "exact sharded" | tee -a gpurun_out/exact_sharded_r2y.log
