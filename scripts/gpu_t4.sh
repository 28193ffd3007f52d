#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_infer.py -q -m gpu -x -k "latency or chain or ddpm or philox or action" 2>&1 | tail -4
for f in 1 0; do echo "== UVA_FUSE_INPROJ=$f"; for B in 56 16; do UVA_FUSE_INPROJ=$f python scripts/lat_breakdown.py $B 2>&1 | grep "lib sample"; done; done
