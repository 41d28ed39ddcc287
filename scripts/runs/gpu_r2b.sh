#!/bin/bash
mkdir -p gpurun_out
CMD="python scripts/wave_probe.py 300000 128 16 64 4096 10000"
$CMD > gpurun_out/wave_probe_r2b.log 2>&1; echo rc=$?; head -2 gpurun_out/wave_probe_r2b.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r2b_wave_launches.csv $CMD > gpurun_out/ncu_r2b.log 2>&1
tail -2 gpurun_out/ncu_r2b.log
