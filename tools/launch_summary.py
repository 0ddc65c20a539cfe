#!/usr/bin/env python
"""Aggregate an ncu launch list (--metrics gpu__time_duration.sum --csv) by kernel: usage launch_summary.py list.csv"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot = collections.OrderedDict()
for r in rows[1:]:
    name = re.sub(r"\(.*", "", r[ki]).replace("void ", "")
    v = float(r[vi].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[ui], 1.0)
    c = tot.setdefault(name, [0, 0.0])
    c[0] += 1
    c[1] += v
s = sum(v for _, v in tot.values())
print("%-84s %6s %12s %7s" % ("kernel", "count", "total_us", "share"))
for k, (n, v) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print("%-84s %6d %12.1f %6.1f%%" % (k[:84], n, v, 100.0 * v / s))
print("%-84s %6d %12.1f" % ("total", sum(n for n, _ in tot.values()), s))
