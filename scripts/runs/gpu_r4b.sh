#!/bin/bash
mkdir -p gpurun_out
for nb in 6 12 24; do
  B200HNSW_PIPE_NBUF=$nb timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2958${nb:0:1} bench.py --gpus 2 > gpurun_out/r4b_nbuf$nb.json 2> gpurun_out/r4b_err_$nb.log || tail -20 gpurun_out/r4b_err_$nb.log
  python scripts/bench_brief.py "N=2 p2p nbuf $nb" < gpurun_out/r4b_nbuf$nb.json
  python -c "
import json; l=json.load(open('gpurun_out/r4b_nbuf$nb.json')); print('   event_kernel_ms', l['roofline']['event_kernel_ms'])"
done
