#!/usr/bin/env python
"""Aggregate `ncu --page source --csv --print-source cuda,sass` output per CUDA source line: samples, instructions.
usage: ncu -i rep --page source --csv --print-source cuda,sass --kernel-name K --launch-skip n --launch-count 1 | python tools/ncu_lines.py [top]"""
import csv
import sys

top = int(sys.argv[1]) if len(sys.argv) > 1 else 30
rows = list(csv.reader(sys.stdin))
cur_file, hdr, agg = "", None, {}
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        hdr = r; continue
    if hdr is None or len(r) < len(hdr) or r[2] != "-":   # keep only the per-line summary rows (Address == "-")
        continue
    try:
        line = int(r[0]); samples = int(r[hdr.index("# Samples")]); inst = int(r[hdr.index("Instructions Executed")])
    except ValueError:
        continue
    a = agg.setdefault((cur_file, line), [0, 0, r[1].strip()[:110]])
    a[0] += samples; a[1] += inst
tot = sum(a[0] for a in agg.values()) or 1
toti = sum(a[1] for a in agg.values()) or 1
print("total samples %d, instructions %d" % (tot, toti))
for (f, l), a in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    print("%5.1f%% smp %5.1f%% inst  %s:%d  %s" % (100.0 * a[0] / tot, 100.0 * a[1] / toti, f, l, a[2]))
