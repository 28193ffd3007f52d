#!/bin/bash
for gmul in 1 2 4 7; do for ov in 0 1; do
  echo "== grid_mult=$gmul overlap=$ov"
  UVA_GRID_MULT=$gmul UVA_OVERLAP=$ov timeout 300 python bench.py --steps 2 --warmup 3 --denoise-steps 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step']/20*1000/4, 'us per 32768-row step', d['value']/1e6, d['clocks'])"
done; done
