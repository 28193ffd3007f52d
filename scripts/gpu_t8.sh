#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_train.py -q -m gpu -x 2>&1 | tail -4
python scripts/train_diag.py 2>&1 | tail -4
echo "== UVA_FINAL_TC=0 UVA_FUSE_COLSUM=0"
UVA_FINAL_TC=0 UVA_FUSE_COLSUM=0 python scripts/train_diag.py 2>&1 | head -1
