#!/bin/bash
# tests + a short sampling bench (20 of 100 denoise steps); extra env settings come from the caller
mkdir -p gpurun_out
echo "== all gpu tests"; timeout 1200 python -m pytest tests -q -m gpu -x 2>&1 | tail -4
for ov in 1 0; do
  echo "== bench overlap=$ov"
  UVA_OVERLAP=$ov timeout 300 python bench.py --steps 2 --warmup 3 --denoise-steps 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step']/20*1000/4, 'us per 32768-row step', d['value']/1e6, 'M row-steps/s', d['clocks'])"
done
CMD="python bench.py --rows 32768 --denoise-steps 2 --steps 1 --warmup 1 --no-cpu-baseline"
export UVA_OVERLAP=0
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_ab.csv $CMD > gpurun_out/ncu_launch.log 2>&1
python scripts/agg_launches.py gpurun_out/launches_ab.csv 8
