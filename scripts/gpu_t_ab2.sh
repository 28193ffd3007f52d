#!/bin/bash
# A/B of two library builds on the C2 training step: build/libuva_base.so vs build/libuva_new.so
timeout 600 python -m pytest tests/test_gpu_train.py -q -x 2>&1 | tail -2
for v in base new base new; do
  UVA_HEAD_LIB=$PWD/build/libuva_$v.so timeout 300 python bench.py --workload train --steps 50 --warmup 10 2>&1 | tail -1 | python -c "
import sys, json; d=json.loads(sys.stdin.read()); x=d['detail']; print('$v', 'ms', round(x['ms_per_step'],3), 'p50', round(x['p50_ms'],3))"
done
