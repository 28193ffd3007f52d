#!/bin/bash
# round-1 checkpoint 9 (v13): GPU tests, smoke, both bench arms (with extras), train + sweep workloads, launch list of one training step
mkdir -p gpurun_out
echo "== gpu tests"; timeout 1500 python -m pytest tests -q -m gpu 2>&1 | tail -4
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
echo "== bench reference arm"; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 2>&1 | tail -1 > gpurun_out/bench_ref_v13.json; cut -c1-200 gpurun_out/bench_ref_v13.json
echo "== bench"; timeout 1200 python bench.py --steps 3 --warmup 3 2>&1 | tail -1 > gpurun_out/bench_v13.json; cut -c1-400 gpurun_out/bench_v13.json; python -c "
import json; d=json.load(open('gpurun_out/bench_v13.json')); print(json.dumps(d.get('extras'))[:1500]); print(d['roofline']['achieved'], d['roofline']['frac'], d['e2e'], d['clocks'])"
echo "== train bench"; timeout 600 python bench.py --workload train --steps 50 --warmup 10 2>&1 | tail -1 > gpurun_out/train_v13.json; cut -c1-600 gpurun_out/train_v13.json
echo "== sweep"; timeout 900 python bench.py --workload sweep 2>&1 | tail -1 > gpurun_out/sweep_v13.json; cut -c1-1200 gpurun_out/sweep_v13.json
echo "== train launch list"; UVA_ACT_OVERLAP=0 UVA_TRAIN_GRAPH=0 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_train_v13.csv python bench.py --workload train --steps 1 --warmup 2 > gpurun_out/ncu_train_v13.log 2>&1; tail -2 gpurun_out/ncu_train_v13.log | cut -c1-300; wc -l gpurun_out/launches_train_v13.csv
