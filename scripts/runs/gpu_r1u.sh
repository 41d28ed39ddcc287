#!/bin/bash
mkdir -p gpurun_out
for v in A B C D; do
  if [ $v = A ]; then unset B200HNSW_LIB; else export B200HNSW_LIB=$PWD/hnsw_b200/lib/exp_$v.so; fi
  echo "== lib $v"
  timeout 600 python scripts/k1_variants.py 1000000 0 10000,100000 2>&1 | grep variant
done > gpurun_out/k1_exp_r1u.log 2>&1
cat gpurun_out/k1_exp_r1u.log
