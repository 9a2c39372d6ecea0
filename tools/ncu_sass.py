#!/usr/bin/env python
"""SASS rows of a source-line range from `ncu --page source --csv --print-source cuda,sass`, with samples and stall reasons.
usage: ncu -i rep --page source --csv --print-source cuda,sass | python tools/ncu_sass.py file.cu 161 215 [min_samples]"""
import csv
import sys

fname, lo, hi = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
mins = int(sys.argv[4]) if len(sys.argv) > 4 else 5
rows = list(csv.reader(sys.stdin))
curfile, hdr, line = "", None, -1
tot_s = tot_i = 0
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        curfile = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        hdr = r; ix = {n: i for i, n in enumerate(hdr)}; continue
    if hdr is None or len(r) < len(hdr):
        continue
    if r[0] != "":
        try:
            line = int(r[0])
        except ValueError:
            line = -1
        continue
    if curfile != fname or not (lo <= line <= hi):
        continue
    try:
        smp = int(r[ix["# Samples"]]); inst = int(r[ix["Instructions Executed"]])
    except ValueError:
        continue
    tot_s += smp; tot_i += inst
    if smp >= mins:
        st = {n[6:]: int(r[ix[n]]) for n in hdr if n.startswith("stall_") and "Not" not in n and r[ix[n]] not in ("0", "")}
        print("%4d %-52s smp %5d inst %9d %s" % (line, r[3].strip()[:52], smp, inst, st))
print("total samples %d instructions %d" % (tot_s, tot_i))
