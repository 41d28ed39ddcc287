#!/bin/bash
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_search_parity_gpu.py tests/test_sharded_gpu.py -m gpu -x -q 2>&1 | tail -6) | tee gpurun_out/r4c_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29591 bench.py --gpus 2 > gpurun_out/r4c_bench_2gpu.json 2> gpurun_out/r4c_err.log || tail -20 gpurun_out/r4c_err.log
python scripts/bench_brief.py "N=2 p2p warp-merge" < gpurun_out/r4c_bench_2gpu.json
python -c "
import json; l=json.load(open('gpurun_out/r4c_bench_2gpu.json')); print('   event_kernel_ms', l['roofline']['event_kernel_ms'])"
