#!/bin/bash
# round 2, call 8v: a few ncu counters of spec_insert_kernel (2 000 insertions into a 1 M-row graph, window 32)
mkdir -p gpurun_out
CMD="python scripts/k2s_probe.py 1000000 128 2000 32"
timeout 30 $CMD > gpurun_out/r8v_plain.log 2>&1 && tail -1 gpurun_out/r8v_plain.log &&
timeout 45 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,lts__t_sector_hit_rate.pct,launch__grid_size,launch__registers_per_thread,launch__shared_mem_per_block_dynamic --clock-control none -k regex:spec_insert -c 1 --csv --log-file gpurun_out/r8v_k2s_counters.csv $CMD > gpurun_out/r8v_ncu.log 2>&1
tail -1 gpurun_out/r8v_ncu.log; tail -9 gpurun_out/r8v_k2s_counters.csv | cut -c1-220
