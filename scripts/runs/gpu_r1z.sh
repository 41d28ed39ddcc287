#!/bin/bash
mkdir -p gpurun_out
CMD="python scripts/wave_probe.py 300000 128 16 64 4096 10000"
$CMD > gpurun_out/wave_probe_r1z.log 2>&1; echo rc=$?; cat gpurun_out/wave_probe_r1z.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r1z_wave_launches.csv $CMD > gpurun_out/ncu_r1z.log 2>&1
tail -2 gpurun_out/ncu_r1z.log
