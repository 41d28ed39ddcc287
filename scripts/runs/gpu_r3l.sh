#!/bin/bash
mkdir -p gpurun_out
run() { # name, env, args
  env $2 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29540 + RANDOM % 50)) bench.py --gpus 2 --no-cpu-baseline $3 > gpurun_out/r3l_$1.json 2>/dev/null
  python - <<PY
import json
l=json.load(open('gpurun_out/r3l_$1.json'))
print('$1', l['value'], l['ms_per_step'], 'e2e', l['e2e']['value'])
PY
}
run prio_hi_d2 B200HNSW_SIDE_PRIORITY=-1 "--inflight 2"
run prio_lo_d2 B200HNSW_SIDE_PRIORITY=0 "--inflight 2"
run prio_hi_d3 B200HNSW_SIDE_PRIORITY=-1 "--inflight 3"
run prio_hi_d4 B200HNSW_SIDE_PRIORITY=-1 "--inflight 4"
