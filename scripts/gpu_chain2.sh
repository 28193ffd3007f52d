#!/bin/bash
export UVA_VERBOSE=1
timeout 300 python -m pytest tests/test_gpu_infer.py -q -m gpu -x -k "chain or latency" 2>&1 | tail -5
for cl in 1 4; do echo "== UVA_CHAIN_CL=$cl"; UVA_CHAIN_CL=$cl timeout 120 python scripts/lat_breakdown.py 1 2>&1 | grep -E "lib sample|full sample|co-resident"; done
