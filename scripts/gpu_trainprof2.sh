#!/bin/bash
# full ncu capture of the training GEMM kinds (video head, 16384 rows)
mkdir -p gpurun_out
export UVA_ACT_OVERLAP=0
CMD="python bench.py --workload train --steps 1 --warmup 1"
$CMD 2>&1 | tail -1 | cut -c1-200
# first forward of the video head: launches 0,1 = cond_embed, AdaLN; then (SILU_SAVE, GATED_SAVE) x 6
timeout 900 ncu --set full --clock-control none --import-source on -k regex:umma_gemm2_kernel -s 2 -c 2 -o gpurun_out/prof_train_fwd $CMD > gpurun_out/ncu_t1.log 2>&1
# video backward: the launches of the second training step; skip the first step (14 + 14 fwd, 44 bwd ... ) by count
timeout 900 ncu --set full --clock-control none --import-source on -k regex:umma_gemm2_kernel -s 56 -c 6 -o gpurun_out/prof_train_bwd $CMD > gpurun_out/ncu_t2.log 2>&1
ls -la gpurun_out/*.ncu-rep
for f in prof_train_fwd prof_train_bwd; do
  ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/$f.csv 2>/dev/null
  python scripts/ncu_summary.py gpurun_out/$f.csv | grep -E "Kernel Name|time_duration|tensor_cycles|dram__bytes|dram_throughput|grid_size|registers" 
done
