#!/bin/bash
# v12 profiles of the final code: sampling launch list, full ncu capture of the sampling GEMM kinds and of the chain kernel
mkdir -p gpurun_out
CMD="python bench.py --rows 65536 --denoise-steps 2 --steps 1 --warmup 1 --no-cpu-baseline --no-extras"
export UVA_OVERLAP=0
$CMD 2>&1 | tail -1 | cut -c1-200
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_v12.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "== launch list done"; python scripts/agg_launches.py gpurun_out/launches_v12.csv 12
timeout 900 ncu --set full --clock-control none --import-source on -k regex:umma_gemm2_kernel -s 26 -c 6 -o gpurun_out/prof_gemm2_v12 $CMD > gpurun_out/ncu_full.log 2>&1
timeout 300 python scripts/lat_breakdown.py 1 2>&1 | tail -4
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_denoise_chain -s 2 -c 1 -o gpurun_out/prof_chain_v12 python scripts/lat_breakdown.py 1 > gpurun_out/ncu_full3.log 2>&1
for f in prof_gemm2_v12 prof_chain_v12; do ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/$f.csv 2>/dev/null; python scripts/ncu_summary.py gpurun_out/$f.csv > gpurun_out/$f.txt; done
grep -E "Kernel Name|time_duration|tensor_cycles|dram__bytes|dram_throughput|issue_active|warps_active" gpurun_out/prof_gemm2_v12.txt gpurun_out/prof_chain_v12.txt | cut -c1-170
ncu -i gpurun_out/prof_chain_v12.ncu-rep --page details --csv 2>/dev/null | grep -i "stall\|Issue Slots\|No Eligible\|Executed Ipc" | cut -c1-220 | head -30 > gpurun_out/prof_chain_v12_details.txt
rm -f gpurun_out/*.ncu-rep
