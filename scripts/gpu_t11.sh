#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_train.py -q -m gpu -x 2>&1 | tail -3
python scripts/train_diag.py 2>&1 | sed -n 1,3p
echo "== UVA_EAGER_BACKWARD=0"; UVA_EAGER_BACKWARD=0 python scripts/train_diag.py 2>&1 | sed -n 3,3p
timeout 300 python bench.py --workload train --steps 20 --warmup 5 2>&1 | tail -1 | cut -c1-330
