#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --rows 32768 --denoise-steps 2 --steps 1 --warmup 1 --no-cpu-baseline"
export UVA_OVERLAP=0 UVA_VERBOSE=1
for q in 1 0; do
export UVA_QUAD=$q
$CMD 2>&1 | grep uva
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_quad$q.csv $CMD > gpurun_out/ncu_launch.log 2>&1
python scripts/agg_launches.py gpurun_out/launches_quad$q.csv 6
done
