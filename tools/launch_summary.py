"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel, launches,
total time and share.  usage: python tools/launch_summary.py LAUNCHES.csv"""
import csv
import re
import sys
from collections import OrderedDict

rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith('==')]
rd = csv.DictReader(lines)
tot = OrderedDict()
for r in rd:
    if r.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    name = re.sub(r'\(.*', '', r['Kernel Name'])[:70]
    v = float(r['Metric Value'].replace(',', ''))
    unit = r.get('Metric Unit', 'ns')
    ms = v * {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3}.get(unit, 1e-6)
    n, t = tot.get(name, (0, 0.0))
    tot[name] = (n + 1, t + ms)
total = sum(t for _, t in tot.values())
for name, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print('%-70s launches %4d  total %10.3f ms  %5.1f%%' % (name, n, t, 100.0 * t / total))
