#!/usr/bin/env python3
"""Per source line: executed warp instructions (per tile) and stall samples, from
`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass`.  Usage: ncu_top_lines.py file.csv [n_tiles] [min_pct]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
ntiles = float(sys.argv[2]) if len(sys.argv) > 2 else 32768
minpct = float(sys.argv[3]) if len(sys.argv) > 3 else 0.4
fname, hdr, out = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname, hdr = r[1].split("/")[-1], None
    elif r[0] == "Line No":
        hdr = r
    elif hdr and r[0].isdigit() and len(r) == len(hdr):
        try:
            out.append((fname, int(r[0]), int(r[hdr.index("Instructions Executed")]), int(r[hdr.index("# Samples")]), r[1]))
        except ValueError:
            pass
tot = sum(o[2] for o in out)
ts = sum(o[3] for o in out)
print("total exec/tile %.0f, samples %d" % (tot / ntiles, ts))
for f, l, e, s, src in out:
    if e > tot * minpct / 100 or s > ts * minpct / 100:
        print(f"{f[:14]:14s} {l:5d} exec/tile {e / ntiles:8.1f} {100 * e / tot:5.2f}%  smp {100 * s / max(ts, 1):5.2f}%  {src.strip()[:90]}")
