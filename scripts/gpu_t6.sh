#!/bin/bash
for B in 56 16 4; do timeout 120 python scripts/lat_breakdown.py $B 2>&1 | grep "lib sample"; done
export UVA_HEAD_LIB=$PWD/build/libuva_exp.so
echo "== early trigger"
for B in 56 16 4; do timeout 120 python scripts/lat_breakdown.py $B 2>&1 | grep "lib sample"; done
timeout 600 python -m pytest tests/test_gpu_infer.py -q -m gpu -x -k "latency or ddpm or philox" 2>&1 | tail -3
