"""Summarises an ncu launch list (gpu__time_duration.sum per launch) per kernel."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ki, vi = hdr.index('Kernel Name'), hdr.index('Metric Value')
agg = collections.OrderedDict()
for r in rows[1:]:
    name = r[ki].split('(')[0].replace('void ', '').replace('mr::', '')
    v = float(r[vi].replace(',', ''))
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"{'kernel':45s} {'n':>5s} {'total us':>10s} {'avg us':>9s} {'share':>6s}")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:45s} {n:5d} {t/1e3:10.1f} {t/n/1e3:9.2f} {t/tot*100:5.1f}%")
