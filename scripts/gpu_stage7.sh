#!/bin/bash
mkdir -p gpurun_out
echo "== infer tests" ; timeout 1200 python -m pytest tests/test_gpu_infer.py -q -m gpu -x -k "not gemm" 2>&1 | tail -8
for f in 1 0; do for ov in 0 1; do
  echo "== bench fuse_ln=$f overlap=$ov"
  UVA_FUSE_LN=$f UVA_OVERLAP=$ov timeout 300 python bench.py --steps 2 --warmup 3 --denoise-steps 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step']/20*1000/4, 'us per 32768-row step', d['value']/1e6, d['roofline']['achieved'], d['clocks'])"
done; done
