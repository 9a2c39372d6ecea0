#!/usr/bin/env python
"""Compact table of the headline metrics of every launch in an ncu report: ncu -i rep --page raw --csv | python tools/ncu_raw.py"""
import csv
import sys

rows = list(csv.reader(sys.stdin))
hdr, units = rows[0], rows[1]
want = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "time"), ("launch__grid_size", "grid"), ("launch__block_size", "blk"),
        ("launch__registers_per_thread", "regs"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"), ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"), ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst"),
        ("l1tex__t_sector_hit_rate.pct", "l1hit%"), ("lts__t_sector_hit_rate.pct", "l2hit%"), ("smsp__inst_executed.sum", "inst")]
idx = [(hdr.index(m), n) for m, n in want if m in hdr]
print(" | ".join(n + ("[" + units[i] + "]" if units[i] else "") for i, n in idx))
for r in rows[2:]:
    out = []
    for i, n in idx:
        v = r[i]
        if n == "kernel":
            v = v.split("(")[0].replace("lego::", "")
        else:
            try:
                v = "%.4g" % float(v.replace(",", ""))
            except ValueError:
                pass
        out.append(v)
    print(" | ".join(out))
