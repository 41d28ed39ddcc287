#!/bin/bash
mkdir -p gpurun_out
# K1 full capture (one 10k-query launch on the 1M x 128 graph), after the plain run exits 0
CMD="python scripts/k1_variants.py 1000000 0 10000"
$CMD > gpurun_out/r3n_k1_plain.log 2>&1 || { tail gpurun_out/r3n_k1_plain.log; exit 1; }
tail -2 gpurun_out/r3n_k1_plain.log
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:traverse_kernel -s 3 -c 1 -f -o gpurun_out/r3n_k1_full $CMD > gpurun_out/r3n_ncu_k1.log 2>&1
tail -2 gpurun_out/r3n_ncu_k1.log
# K3 GEMM full capture
CMD2="python scripts/exact_probe.py 1000000,128,10000,10,0"
$CMD2 > gpurun_out/r3n_exact_plain.log 2>&1 || { tail gpurun_out/r3n_exact_plain.log; exit 1; }
tail -3 gpurun_out/r3n_exact_plain.log
ncu --set full --clock-control none --import-source on -k regex:exact_gemm_kernel --launch-skip 1 --launch-count 1 -f -o gpurun_out/r3n_exact_gemm_full $CMD2 > gpurun_out/r3n_ncu_exact.log 2>&1
tail -2 gpurun_out/r3n_ncu_exact.log
# launch list of the bench command
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r3n_bench_short.json 2>/dev/null; echo "bench rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/r3n_bench_launches.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r3n_ncu_bench.log 2>&1
tail -2 gpurun_out/r3n_ncu_bench.log
ls -la gpurun_out/ | grep r3n
