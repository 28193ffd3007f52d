#!/bin/bash
mkdir -p gpurun_out
echo "== all gpu tests" ; timeout 1200 python -m pytest tests -q -m gpu -x 2>&1 | tail -8
for pr in 1 0; do
  echo "== bench pair=$pr"
  UVA_PAIR=$pr timeout 300 python bench.py --steps 2 --warmup 3 --denoise-steps 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, 'M row-steps/s', d['roofline']['achieved'], d['clocks'])"
  UVA_OVERLAP=0 UVA_PAIR=$pr timeout 300 python bench.py --steps 2 --warmup 3 --denoise-steps 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('no overlap', d['value']/1e6, 'M row-steps/s', d['roofline']['achieved'], d['clocks'])"
done
