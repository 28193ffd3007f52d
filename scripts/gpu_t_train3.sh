#!/bin/bash
mkdir -p gpurun_out
for i in 1 2 3; do
  timeout 300 python bench.py --workload train --steps 20 --warmup 5 2>&1 | tail -1 | python -c "
import sys, json; d=json.loads(sys.stdin.read()); x=d['detail']; print('run', $i, 'ms', x['ms_per_step'], 'p50', x['p50_ms'], 'max', x['max_ms'])"
done
timeout 300 python bench.py --workload train --steps 100 --warmup 10 2>&1 | tail -1 | python -c "
import sys, json; d=json.loads(sys.stdin.read()); x=d['detail']; print('run 100', 'ms', x['ms_per_step'], 'p50', x['p50_ms'], 'max', x['max_ms'])"
nvidia-smi --query-gpu=clocks.sm,power.draw,temperature.gpu,clocks_throttle_reasons.active --format=csv
