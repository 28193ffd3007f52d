#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --rows 32768 --denoise-steps 2 --steps 1 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/prof_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1a.csv $CMD > gpurun_out/ncu_launch.log 2>&1
tail -2 gpurun_out/prof_plain.log
