#!/bin/bash
# timing diagnosis: GEMMs only / row kernels only / both, with and without the two-lane overlap
for sk in 0 1 2; do for ov in 0 1; do
  echo "== skip=$sk overlap=$ov"
  UVA_DEBUG_SKIP=$sk UVA_OVERLAP=$ov timeout 300 python bench.py --steps 2 --warmup 3 --denoise-steps 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step']/20*1000/4, 'us per 32768-row step', d['clocks'])"
done; done
