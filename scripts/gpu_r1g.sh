#!/bin/bash
# round-1 checkpoint 3 (v5): GPU tests, smoke, both bench arms, training bench + training launch list
mkdir -p gpurun_out
echo "== gpu tests"; timeout 1500 python -m pytest tests -q -m gpu 2>&1 | tail -6
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
echo "== bench reference arm"; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 2>&1 | tail -1 > gpurun_out/bench_ref_v5.json; cut -c1-200 gpurun_out/bench_ref_v5.json
echo "== bench"; timeout 1200 python bench.py --steps 3 --warmup 3 2>&1 | tail -1 > gpurun_out/bench_v5.json; cut -c1-3000 gpurun_out/bench_v5.json
echo "== train bench"; timeout 600 python bench.py --workload train --steps 20 --warmup 5 2>&1 | tail -1 > gpurun_out/train_v5.json; cut -c1-1500 gpurun_out/train_v5.json
CMD="python bench.py --workload train --steps 1 --warmup 1"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_train_v5.csv $CMD > gpurun_out/ncu_train.log 2>&1
echo "== train launch list done"
python - <<'PY'
import collections, csv, re
lines = [l for l in open('gpurun_out/launches_train_v5.csv') if not l.startswith('==')]
rows = list(csv.DictReader(lines))
half = rows[len(rows)//2:]
agg = collections.OrderedDict()
for d in half:
    name = re.sub(r'\(.*', '', d['Kernel Name'])[:110]
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += float(d['Metric Value'].replace(',', ''))
tot = sum(a[1] for a in agg.values())
print('second half of the launch list: total %.1f us, %d launches' % (tot/1e3, len(half)))
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
    print(f"{t/1e3:10.1f} us {n:4d}x {t/n/1e3:9.1f} us avg {100*t/tot:5.1f}%  {k}")
PY
