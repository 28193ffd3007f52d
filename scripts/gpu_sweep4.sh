#!/bin/bash
for c in 9472 18944 37888 65536; do for ov in 0 1; do
  echo "== chunk $c overlap $ov"
  UVA_CHUNK_ROWS=$c UVA_OVERLAP=$ov timeout 300 python bench.py --steps 2 --warmup 3 --denoise-steps 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, 'M row-steps/s', d['clocks'])"
done; done
