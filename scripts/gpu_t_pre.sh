#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_infer.py -q -x -k "prephase or latency_path" 2>&1 | tail -15
for v in 0 1 0 1; do
  UVA_FUSE_LNPRE=$v timeout 300 python bench.py --workload latency 2>&1 | tail -1 | python -c "
import sys, json; d=json.loads(sys.stdin.read())['detail']; print('lnpre', $v, {k: (round(x['p50_ms'],3), x['launches_per_call']) for k, x in d.items()})"
done
