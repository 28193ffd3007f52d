#!/bin/bash
mkdir -p gpurun_out
echo "== train tests"; timeout 1200 python -m pytest tests/test_gpu_train.py -q -m gpu -x 2>&1 | tail -4
echo "== train bench"; timeout 600 python bench.py --workload train --steps 10 --warmup 3 2>&1 | tail -1 | cut -c1-700
