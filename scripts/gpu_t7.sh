#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_infer.py -q -m gpu -x -k "noise_graph or action" 2>&1 | tail -6
timeout 300 python bench.py --workload latency 2>&1 | tail -1 | cut -c1-900
