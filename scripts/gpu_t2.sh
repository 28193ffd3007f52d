#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | tail -4
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
timeout 600 python bench.py --workload train --steps 20 --warmup 5 2>&1 | tail -1 | cut -c1-700
