#!/bin/bash
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q 2>&1 | tail -6) | tee gpurun_out/r4d_pytest.log
for fl in memop kernel; do
  B200HNSW_PEER_FLAGS=$fl timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29593 bench.py --gpus 2 > gpurun_out/r4d_bench_2gpu_$fl.json 2> gpurun_out/r4d_err_$fl.log || tail -20 gpurun_out/r4d_err_$fl.log
  python scripts/bench_brief.py "N=2 p2p flags=$fl" < gpurun_out/r4d_bench_2gpu_$fl.json
  python -c "
import json; l=json.load(open('gpurun_out/r4d_bench_2gpu_$fl.json')); print('   event_kernel_ms', l['roofline']['event_kernel_ms'])"
done
