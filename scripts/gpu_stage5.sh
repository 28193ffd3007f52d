#!/bin/bash
mkdir -p gpurun_out
echo "== pair gemm" ; timeout 600 python -m pytest tests/test_gpu_infer.py -q -m gpu -x -k "pair" 2>&1 | tail -30
