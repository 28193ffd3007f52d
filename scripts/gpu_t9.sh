#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_optim.py -q -m gpu -x 2>&1 | tail -4
python scripts/train_diag.py 2>&1 | tail -4
echo "== UVA_TRAIN_GRAPH=0"
UVA_TRAIN_GRAPH=0 python scripts/train_diag.py 2>&1 | sed -n 2,3p
timeout 300 python bench.py --workload train --steps 20 --warmup 5 2>&1 | tail -1 | cut -c1-600
