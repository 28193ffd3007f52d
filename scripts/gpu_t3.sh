#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_infer.py -q -m gpu -x 2>&1 | tail -6
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras 2>&1 | tail -1 | cut -c1-330
python scripts/lat_breakdown.py 56 2>&1 | grep "lib sample"
