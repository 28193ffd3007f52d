#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_infer.py -q -m gpu -x 2>&1 | tail -8
