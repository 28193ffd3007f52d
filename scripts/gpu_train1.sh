#!/bin/bash
mkdir -p gpurun_out
echo "== train tests" ; timeout 1200 python -m pytest tests/test_gpu_train.py -q -m gpu -x 2>&1 | tail -40
