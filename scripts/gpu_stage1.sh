#!/bin/bash
# staged first GPU run: fp32 path, then the tcgen05 GEMM core, then everything, then a short bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
echo "== stage a: fp32 path" ; timeout 300 python -m pytest tests/test_gpu_infer.py -q -m gpu -k "small_geometry" 2>&1 | tail -15
echo "== stage b: umma gemm" ; timeout 300 python -m pytest tests/test_gpu_infer.py -q -m gpu -k "umma_gemm" 2>&1 | tail -30
echo "== stage c: all" ; timeout 900 python -m pytest tests/test_gpu_infer.py -q -m gpu 2>&1 | tail -40
echo "== stage d: smoke" ; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -5
echo "== stage e: bench" ; timeout 600 python bench.py --steps 2 --warmup 3 2>&1 | tail -3
