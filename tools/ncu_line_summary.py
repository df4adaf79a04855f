#!/usr/bin/env python3
"""Per source line totals from `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass`:
executed warp instructions and stall samples per file:line.  Usage: ncu_line_summary.py file.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 45
fname, hdr, out = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
        hdr = None
    elif r[0] == "Function Name":
        continue
    elif r[0] == "Line No":
        hdr = r
    elif hdr and r[0].isdigit() and len(r) == len(hdr):
        iE, iS = hdr.index("Instructions Executed"), hdr.index("# Samples")
        iB, iN, iL = hdr.index("stall_barrier"), hdr.index("stall_no_inst"), hdr.index("stall_long_sb")
        try:
            out.append((fname, int(r[0]), r[1].strip()[:90], int(r[iE]), int(r[iS]), int(r[iB] or 0), int(r[iN] or 0), int(r[iL] or 0)))
        except ValueError:
            pass
totE = sum(o[3] for o in out)
totS = sum(o[4] for o in out)
print("total executed", totE, "samples", totS)
print("%-18s %5s %7s %7s %6s %6s %6s  %s" % ("file", "line", "exec%", "smpl%", "barr", "noinst", "longsb", "source"))
for o in sorted(out, key=lambda o: -o[4])[:top]:
    print("%-18s %5d %6.2f%% %6.2f%% %6d %6d %6d  %s" % (o[0], o[1], 100 * o[3] / totE, 100 * o[4] / max(totS, 1), o[5], o[6], o[7], o[2]))
