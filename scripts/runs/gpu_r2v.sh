#!/bin/bash
mkdir -p gpurun_out
echo "== C3 shape: 500k x 768 cosine M=32"; timeout 900 python scripts/wave_probe.py 500000 768 32 64 4096 10000 cosine 2>&1 | tail -6
echo "== C4 shard shape: 1.25M x 96 L2 M=16"; timeout 900 python scripts/wave_probe.py 1250000 96 16 64 4096 10000 l2 2>&1 | tail -6
