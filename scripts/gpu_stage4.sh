#!/bin/bash
# MN-major / split-K GEMM layouts + regression of the inference tests
mkdir -p gpurun_out
echo "== gemm layouts" ; timeout 600 python -m pytest tests/test_gpu_infer.py -q -m gpu -k "umma_gemm" 2>&1 | tail -30
echo "== all infer" ; timeout 900 python -m pytest tests/test_gpu_infer.py -q -m gpu -k "not umma_gemm" 2>&1 | tail -15
