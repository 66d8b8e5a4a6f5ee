#!/usr/bin/env python3
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time and share per kernel name.
usage: launch_summary.py launches.csv [skip_first_n_launches]"""
import csv
import sys
from collections import OrderedDict

lines = [l for l in open(sys.argv[1]) if l.startswith('"')]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = list(csv.DictReader(lines))
agg = OrderedDict()
tot = 0.0
for r in rows:
    if r["Metric Name"] != "gpu__time_duration.sum" or int(r["ID"]) < skip:
        continue
    ns = float(r["Metric Value"].replace(",", ""))
    if r["Metric Unit"] == "us":
        ns *= 1e3
    elif r["Metric Unit"] == "ms":
        ns *= 1e6
    k = r["Kernel Name"].split("(")[0]
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1
    a[1] += ns
    tot += ns
print("%-40s %8s %12s %7s" % ("kernel", "launches", "total ms", "share"))
for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-40s %8d %12.3f %6.1f%%" % (k, n, ns / 1e6, 100 * ns / tot))
print("%-40s %8d %12.3f" % ("total", sum(a[0] for a in agg.values()), tot / 1e6))
