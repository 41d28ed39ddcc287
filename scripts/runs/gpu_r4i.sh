#!/bin/bash
mkdir -p gpurun_out
for hook in none 1; do
  B200HNSW_TEST_P2P_FAIL=$hook timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29597 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/r4i_$hook.json 2> gpurun_out/r4i_$hook.err || tail -20 gpurun_out/r4i_$hook.err
  python scripts/bench_brief.py "N=2 fail-hook=$hook" < gpurun_out/r4i_$hook.json
  grep "bench.py:" gpurun_out/r4i_$hook.err | head -2
  python -c "
import json; print('   layout:', json.load(open('gpurun_out/r4i_$hook.json'))['config']['layout'][100:230])"
done
