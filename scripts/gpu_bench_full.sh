#!/bin/bash
mkdir -p gpurun_out
echo "== bench reference arm"; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 2>&1 | tail -1 | tee gpurun_out/bench_ref.json
echo "== bench"; timeout 1200 python bench.py --steps 3 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench.json
echo "== sweep"; timeout 900 python bench.py --workload sweep 2>&1 | tail -1 | tee gpurun_out/bench_sweep.json
