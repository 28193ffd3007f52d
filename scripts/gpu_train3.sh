#!/bin/bash
# side-stream overlap of the action head: test, then training bench with and without it
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_train.py -q -m gpu 2>&1 | tail -5
echo "== overlap on";  timeout 300 python bench.py --workload train --steps 20 --warmup 5 2>&1 | tail -1 | cut -c1-700
echo "== overlap off"; UVA_ACT_OVERLAP=0 timeout 300 python bench.py --workload train --steps 20 --warmup 5 2>&1 | tail -1 | cut -c1-700
