#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_infer.py -q -m gpu -x 2>&1 | tail -3
CMD="python bench.py --rows 65536 --denoise-steps 2 --steps 1 --warmup 1 --no-cpu-baseline --no-extras"
UVA_OVERLAP=0 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_v8.csv $CMD > gpurun_out/ncu_launch.log 2>&1
python scripts/agg_launches.py gpurun_out/launches_v8.csv 8
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras 2>&1 | tail -1 | cut -c1-200
