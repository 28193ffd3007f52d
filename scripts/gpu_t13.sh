#!/bin/bash
for a in 1 0; do echo "== autocast $a"; UVA_BENCH_AUTOCAST=$a timeout 300 python bench.py --workload train --steps 20 --warmup 5 2>&1 | tail -1 | cut -c1-330; done
