#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_infer.py -q -k "prenet_graph or module_sample or noise_graph" 2>&1 | tail -30
timeout 300 python scripts/lat_breakdown.py 1 2>&1 | tail -5
timeout 300 python bench.py --workload latency 2>&1 | tail -1 | cut -c1-900
