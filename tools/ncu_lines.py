#!/usr/bin/env python
"""Top source lines of an `ncu --page source --csv --print-source sass,cuda` export, by executed
warp instructions and stall samples.  usage: ncu_lines.py src.csv [N]"""
import csv
import sys

path, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows, cur = [], None
for r in csv.reader(open(path)):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r[0] in ("Function Name", "Line No") or len(r) < 9:
        continue
    if r[2] == "-":  # source-level aggregate row
        try:
            rows.append((cur, int(r[0]), r[1].strip(), int(r[7]), int(r[6])))
        except ValueError:
            pass
ti = sum(x[3] for x in rows) or 1
ts = sum(x[4] for x in rows) or 1
print("total warp-instructions %d, samples %d" % (ti, ts))
for f, ln, src, ins, smp in sorted(rows, key=lambda x: -x[3])[:top]:
    print("%5.1f%% inst %5.1f%% smp  %s:%d  %s" % (100.0 * ins / ti, 100.0 * smp / ts, f, ln, src[:90]))
