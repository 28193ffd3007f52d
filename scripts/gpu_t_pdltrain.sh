#!/bin/bash
UVA_PDL_TRAIN=1 timeout 900 python -m pytest tests/test_gpu_train.py -q -x 2>&1 | tail -2
for v in 0 1 0 1; do
UVA_PDL_TRAIN=$v timeout 300 python bench.py --workload train --steps 50 --warmup 10 2>&1 | tail -1 | python -c "
import sys, json; d=json.loads(sys.stdin.read()); x=d['detail']; print('pdl_train', $v, 'ms', round(x['ms_per_step'],3), 'p50', round(x['p50_ms'],3))"
done
timeout 300 python scripts/sweep_small.py 2304,4096,8192 2>&1 | tail -1
