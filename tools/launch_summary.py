"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.  usage: launch_summary.py csv"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if 'Kernel Name' in r)
hdr = rows[h]
ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
agg, n = collections.OrderedDict(), 0
for r in rows[h + 1:]:
    if len(r) <= vi:
        continue
    v = float(r[vi].replace(',', '')) * {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3}.get(r[ui], 1e-6)
    a = agg.setdefault(r[ki].split('(')[0], [0, 0.0, 0.0])
    a[0] += 1
    a[1] += v
    a[2] = max(a[2], v)
    n += 1
tot = sum(a[1] for a in agg.values())
print(f"{n} launches, {tot:.3f} ms of kernel time")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:28s} launches={a[0]:4d} sum={a[1]:9.3f} ms share={100 * a[1] / tot:5.1f}% longest={a[2]:.3f} ms")
