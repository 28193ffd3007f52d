#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_infer.py -q -x 2>&1 | tail -3
for i in 1 2; do
timeout 300 python bench.py --workload latency 2>&1 | tail -1 | python -c "
import sys, json; d=json.loads(sys.stdin.read())['detail']; print({k: (round(x['p50_ms'],3), round(x['p50_ms_device_noise'],3)) for k, x in d.items()})"
done
timeout 300 python scripts/lat_breakdown.py 1 2>&1 | tail -4
