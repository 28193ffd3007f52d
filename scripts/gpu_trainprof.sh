#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --workload train --steps 1 --warmup 1"
$CMD 2>&1 | tail -1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_train.csv $CMD > gpurun_out/ncu_train.log 2>&1
python - <<'PY'
import collections, csv, re
lines = [l for l in open('gpurun_out/launches_train.csv') if not l.startswith('==')]
rows = list(csv.DictReader(lines))
# keep the last training step only: find last k_qsample with 16384-row video head... simply take the second half
half = rows[len(rows)//2:]
agg = collections.OrderedDict()
for d in half:
    name = re.sub(r'\(.*', '', d['Kernel Name'])[:90]
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += float(d['Metric Value'].replace(',', ''))
tot = sum(a[1] for a in agg.values())
print('second half of the launch list: total %.1f us, %d launches' % (tot/1e3, len(half)))
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:28]:
    print(f"{t/1e3:10.1f} us {n:4d}x {t/n/1e3:9.1f} us avg {100*t/tot:5.1f}%  {k}")
PY
