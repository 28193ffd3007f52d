#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --rows 32768 --denoise-steps 2 --steps 1 --warmup 1 --no-cpu-baseline"
export UVA_OVERLAP=0
$CMD > gpurun_out/prof_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1d.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "== launch list done"
$CMD > gpurun_out/prof_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:umma_gemm2_kernel -s 24 -c 4 -o gpurun_out/prof_gemm2_r1d $CMD > gpurun_out/ncu_full.log 2>&1
echo "== ncu done"; ls -la gpurun_out/*.ncu-rep
