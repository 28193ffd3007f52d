#!/bin/bash
# round-1 checkpoint 2: all GPU tests, smoke, bench (both arms), launch list + full ncu capture of the top kernel
mkdir -p gpurun_out
echo "== gpu tests"; timeout 1500 python -m pytest tests -q -m gpu 2>&1 | tail -4
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
echo "== bench reference arm"; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 2>&1 | tail -1 > gpurun_out/bench_ref_r1f.json; cut -c1-200 gpurun_out/bench_ref_r1f.json
echo "== bench"; timeout 1200 python bench.py --steps 3 --warmup 3 2>&1 | tail -1 > gpurun_out/bench_r1f.json; cut -c1-400 gpurun_out/bench_r1f.json
CMD="python bench.py --rows 65536 --denoise-steps 2 --steps 1 --warmup 1 --no-cpu-baseline --no-extras"
export UVA_OVERLAP=0
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r1f.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "== launch list done"; python scripts/agg_launches.py gpurun_out/launches_r1f.csv 10
timeout 900 ncu --set full --clock-control none --import-source on -k regex:umma_gemm2_kernel -s 26 -c 6 -o gpurun_out/prof_gemm2_r1f $CMD > gpurun_out/ncu_full.log 2>&1
echo "== ncu done"; ls -la gpurun_out/*.ncu-rep
