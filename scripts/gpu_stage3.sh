#!/bin/bash
mkdir -p gpurun_out
echo "== tests" ; timeout 900 python -m pytest tests/test_gpu_infer.py -q -m gpu 2>&1 | tail -15
echo "== bench overlap" ; timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>&1 | tail -1
echo "== bench no overlap" ; UVA_OVERLAP=0 timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>&1 | tail -1
CMD="python bench.py --rows 32768 --denoise-steps 2 --steps 1 --warmup 1 --no-cpu-baseline"
UVA_OVERLAP=0 $CMD > gpurun_out/prof_plain.log 2>&1 && \
UVA_OVERLAP=0 ncu --set full --clock-control none --import-source on -k regex:umma_gemm_tn_kernel -s 1 -c 3 -o gpurun_out/prof_gemm_r1c $CMD > gpurun_out/ncu_full.log 2>&1
echo "== ncu done"; ls -la gpurun_out/*.ncu-rep
