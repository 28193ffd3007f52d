#!/bin/bash
# round-1 checkpoint run: GPU tests, smoke, both bench arms, ncu launch list + full capture of the top kernel
mkdir -p gpurun_out
echo "== gpu tests"; timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | tail -6
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -4
echo "== bench reference arm"; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 2>&1 | tail -1 | tee gpurun_out/bench_ref_r1e.json
echo "== bench"; timeout 900 python bench.py --steps 3 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench_r1e.json
CMD="python bench.py --rows 32768 --denoise-steps 2 --steps 1 --warmup 1 --no-cpu-baseline"
export UVA_OVERLAP=0
$CMD > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1e.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "== launch list done"
$CMD > gpurun_out/prof_plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:umma_gemm2_kernel -s 24 -c 6 -o gpurun_out/prof_gemm2_r1e $CMD > gpurun_out/ncu_full.log 2>&1
echo "== ncu done"; ls -la gpurun_out/*.ncu-rep
