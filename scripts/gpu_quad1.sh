#!/bin/bash
mkdir -p gpurun_out
echo "== gemm tests"; timeout 600 python -m pytest tests/test_gpu_infer.py -q -m gpu -x -k "gemm" 2>&1 | tail -8
echo "== all gpu tests"; timeout 1200 python -m pytest tests -q -m gpu -x 2>&1 | tail -4
for q in 1 0; do for ov in 1 0; do
  echo "== bench quad=$q overlap=$ov"
  UVA_QUAD=$q UVA_OVERLAP=$ov timeout 300 python bench.py --steps 2 --warmup 3 --denoise-steps 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step']/20*1000/4, 'us per 32768-row step', d['value']/1e6, 'M row-steps/s', d['clocks'])"
done; done
