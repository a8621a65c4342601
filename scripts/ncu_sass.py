#!/usr/bin/env python
"""Per-SASS-instruction view of one captured kernel: executed count and stall samples, hot loop only.
usage: ncu_sass.py report.ncu-rep [min_fraction_of_max_exec]"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(l for l in out.splitlines() if l.startswith('"')))
hdr = rows[1]
ia, isrc, iex, ismp = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
body = [r for r in rows[2:] if len(r) > iex and r[iex].isdigit()]
mx = max(int(r[iex]) for r in body)
frac = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
tot_s = sum(int(r[ismp]) for r in body)
print("max exec %d, total samples %d" % (mx, tot_s))
for n, r in enumerate(body):
    e = int(r[iex])
    if e >= frac * mx:
        print("%5d %10d %6d  %s" % (n, e, int(r[ismp]), r[isrc].strip()))
