#!/bin/bash
mkdir -p gpurun_out
echo "== tests" ; timeout 900 python -m pytest tests/test_gpu_infer.py -q -m gpu 2>&1 | tail -25
echo "== bench" ; timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>&1 | tail -1
CMD="python bench.py --rows 32768 --denoise-steps 2 --steps 1 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/prof_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1b.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "== launch list done"
