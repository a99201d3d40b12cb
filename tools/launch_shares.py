#!/usr/bin/env python
"""Per-kernel totals of an ncu `--metrics gpu__time_duration.sum --csv` launch list.  usage: launch_shares.py <launches.csv>"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
H = rows[hdr]
ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[hdr + 1:]:
    name = r[ki].split("(")[0][:70]
    v = float(r[vi].replace(",", ""))
    v = v / 1e3 if r[ui] == "ns" else v * 1e3 if r[ui] == "ms" else v
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print("%-72s %6s %12s %7s" % ("kernel", "count", "total us", "share"))
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-72s %6d %12.1f %6.1f%%" % (n, c, t, 100 * t / tot))
