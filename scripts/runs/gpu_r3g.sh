#!/bin/bash
# round-1 re-entry: full GPU test pass, default bench (e2e with 2 concurrent callers), reference arm, C3 / C4 shapes at full size
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
(timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/r3g_pytest_gpu.log; tail -3 gpurun_out/r3g_pytest_gpu.log
timeout 900 python bench.py > gpurun_out/r3g_bench_1gpu.json 2> gpurun_out/r3g_bench_1gpu.err || tail -20 gpurun_out/r3g_bench_1gpu.err
cut -c1-1800 gpurun_out/r3g_bench_1gpu.json
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r3g_bench_reference_arm.json 2> gpurun_out/r3g_ref.err || tail -5 gpurun_out/r3g_ref.err
cut -c1-300 gpurun_out/r3g_bench_reference_arm.json
echo "== C3: 1M x 768 cosine M=32"
timeout 1200 python scripts/config_probe.py 1000000 768 32 64 cosine 10000 20 1000 > gpurun_out/r3g_c3_1m_768_cosine.json 2> gpurun_out/r3g_c3.err || tail -20 gpurun_out/r3g_c3.err
cat gpurun_out/r3g_c3_1m_768_cosine.json
echo "== C4 (all rows on one GPU): 10M x 96 L2 M=16"
timeout 1200 python scripts/config_probe.py 10000000 96 16 64 l2 10000 20 1000 > gpurun_out/r3g_c4_10m_96_l2.json 2> gpurun_out/r3g_c4.err || tail -20 gpurun_out/r3g_c4.err
cat gpurun_out/r3g_c4_10m_96_l2.json
