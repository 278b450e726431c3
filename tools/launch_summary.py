"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import csv
import sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr, data = None, []
for r in rows:
    if r[0] == 'ID':
        hdr = r
        continue
    if hdr:
        data.append(dict(zip(hdr, r)))
agg = defaultdict(lambda: [0, 0.0])
for d in data:
    name = d['Kernel Name'].split('(')[0]
    v = float(d['Metric Value'].replace(',', ''))
    unit = d['Metric Unit']
    v = v / 1e3 if unit == 'ns' else (v * 1e3 if unit == 'ms' else v)
    agg[name][0] += 1
    agg[name][1] += v
tot = sum(v[1] for v in agg.values())
print(f"{'kernel':60s} {'launches':>8s} {'total_us':>10s} {'share':>6s}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:60]:60s} {v[0]:8d} {v[1]:10.1f} {v[1] / tot * 100:5.1f}%")
