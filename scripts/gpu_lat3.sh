#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_infer.py -q -m gpu -x -k "latency or chain or ragged or action" 2>&1 | tail -4
python scripts/lat_breakdown.py 56 2>&1 | tail -4
python scripts/lat_breakdown.py 16 2>&1 | tail -2
python scripts/lat_breakdown.py 1 2>&1 | tail -2
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 2300 -c 120 --csv --log-file gpurun_out/launches_lat56.csv python scripts/lat56_ncu.py > gpurun_out/ncu_lat56.log 2>&1
python scripts/agg_launches.py gpurun_out/launches_lat56.csv 20
