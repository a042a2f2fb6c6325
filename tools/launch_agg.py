#!/usr/bin/env python
"""Aggregate an ncu `--metrics gpu__time_duration.sum --csv` launch list per kernel: tools/launch_agg.py file.csv [divide_by_steps]"""
import csv, collections, re, sys
rows = list(csv.reader(open(sys.argv[1])))
div = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
hi = [i for i, r in enumerate(rows) if 'Kernel Name' in r][0]
h = rows[hi]; kn = h.index('Kernel Name'); mv = h.index('Metric Value'); mu = h.index('Metric Unit')
agg = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) <= mv: continue
    name = re.sub(r'\(.*', '', r[kn]); v = float(r[mv].replace(',', ''))
    if r[mu] == 'ns': v /= 1000
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-52s %6.1f launches %9.1f us %5.1f%%" % (k[:52], c / div, t / div, 100 * t / tot))
print("total %.1f us" % (tot / div))
