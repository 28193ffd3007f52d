#!/bin/bash
# A/B of an experimental library build (build/libuva_*.so via UVA_HEAD_LIB) against the in-tree one on the C2 step
LIBX=${1:-build/libuva_lb2.so}
for v in base exp base exp; do
  if [ $v = exp ]; then export UVA_HEAD_LIB=$PWD/$LIBX; else unset UVA_HEAD_LIB; fi
  timeout 300 python bench.py --workload train --steps 50 --warmup 10 2>&1 | tail -1 | python -c "
import sys, json; d=json.loads(sys.stdin.read()); x=d['detail']; print('$v', 'ms', x['ms_per_step'], 'p50', x['p50_ms'], 'max', x['max_ms'])"
done
