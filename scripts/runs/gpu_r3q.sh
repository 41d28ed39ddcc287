#!/bin/bash
mkdir -p gpurun_out
for per in 0.02 0.02 0.02 0.02; do
  B200_BENCH_E2E_SAMPLE_PERIOD=$per timeout 600 python bench.py --no-exact --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
l=json.loads(sys.stdin.read()); print('period $per', 'value %.1fM' % (l['value']/1e6), 'e2e %.1fM' % (l['e2e']['value']/1e6), 'single %.1fM' % (l['e2e']['single_caller_value']/1e6), 'samples', l['clocks']['samples'])"
done | tee gpurun_out/r3q_e2e_sampler_ab.log
