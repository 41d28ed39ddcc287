#!/bin/bash
# round 2, call 8h: where the committer's time goes (cycles per node: wait / check / run again / commit)
mkdir -p gpurun_out
B200HNSW_TRACE_ADD=1 timeout 900 python scripts/k2s_probe.py 1000000 128 8000 8,32,64 > gpurun_out/r8h_k2s_1m.log 2>&1
cat gpurun_out/r8h_k2s_1m.log
