#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel: launches, total us, share.
usage: python tools/ncu_launch_summary.py launches.csv > summary.csv"""
import csv
import re
import sys

agg = {}
with open(sys.argv[1]) as f:
    rows = [r for r in csv.reader(l for l in f if l.startswith('"'))]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
for r in rows[1:]:
    name = re.sub(r"\(.*", "", r[ki]).replace("lego::", "")
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1; a[1] += float(r[vi].replace(",", "")) / 1e3
tot = sum(a[1] for a in agg.values()) or 1.0
print("kernel,launches,total_us,share_pct")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%s,%d,%.1f,%.2f" % (k, a[0], a[1], 100.0 * a[1] / tot))
