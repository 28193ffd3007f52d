#!/bin/bash
mkdir -p gpurun_out
echo "== infer tests"; timeout 1200 python -m pytest tests/test_gpu_infer.py -q -m gpu -x -k "not gemm" 2>&1 | tail -6
echo "== latency"; timeout 600 python bench.py --workload latency 2>&1 | tail -1
echo "== latency no chain"; UVA_CHAIN=0 timeout 600 python bench.py --workload latency 2>&1 | tail -1
