#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_train.py -q -m gpu -x 2>&1 | tail -3
for r in 16 0 8 24 32; do echo "== reserve $r"; UVA_TRAIN_RESERVE_SMS=$r python scripts/train_diag.py 2>&1 | sed -n 1,3p; done
