#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_train.py -q -x 2>&1 | tail -3
for v in 0 1 0 1; do
  UVA_EPI_COLSUM=$v timeout 300 python bench.py --workload train --steps 50 --warmup 10 2>&1 | tail -1 | python -c "
import sys, json; d=json.loads(sys.stdin.read()); x=d['detail']; print('epi_colsum', $v, 'ms', x['ms_per_step'], 'p50', x['p50_ms'], 'max', x['max_ms'])"
done
