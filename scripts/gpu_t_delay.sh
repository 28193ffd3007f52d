#!/bin/bash
for v in 384 512 640 0 448 576; do
UVA_CHAIN_HOLD=$v timeout 300 python bench.py --workload latency 2>&1 | tail -1 | python -c "
import sys, json; d=json.loads(sys.stdin.read())['detail']; print('hold', $v, {k: (round(x['p50_ms'],3), round(x['p50_ms_device_noise'],3)) for k, x in d.items()})"
done
timeout 600 python -m pytest tests/test_gpu_infer.py -q -x -k "latency_path or action_head or cluster or other_channel" 2>&1 | tail -2
