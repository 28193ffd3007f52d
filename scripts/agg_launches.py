"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel."""
import collections, csv, re, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
agg = collections.OrderedDict()
for d in csv.DictReader(lines):
    name = re.sub(r'\(.*', '', d['Kernel Name'])[:80]
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += float(d['Metric Value'].replace(',', ''))
tot = sum(a[1] for a in agg.values())
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 14]:
    print(f"{t/1e3:10.1f} us {n:4d}x {t/n/1e3:9.1f} us avg {100*t/tot:5.1f}%  {k}")
