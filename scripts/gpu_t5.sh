#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_infer.py -q -m gpu -x 2>&1 | tail -4
for f in 1 0; do echo "== UVA_PDL=$f"; for B in 56 16 4; do UVA_PDL=$f timeout 120 python scripts/lat_breakdown.py $B 2>&1 | grep "lib sample"; done; done
