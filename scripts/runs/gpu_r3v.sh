#!/bin/bash
mkdir -p gpurun_out
B200HNSW_PIPE_PROFILE=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29562 bench.py --gpus 2 > gpurun_out/r3v_bench_2gpu_prof.json 2>/dev/null
python -c "
import json; l=json.load(open('gpurun_out/r3v_bench_2gpu_prof.json')); print(l['value'], l['ms_per_step'], l.get('pipeline_profile'), l['roofline']['kernel_ms'])"
