#!/usr/bin/env python3
"""Summarise `ncu -i X.ncu-rep --page source --csv` (SASS view): executed warp instructions by
opcode, stall-sample totals, and the hottest instructions.  Usage: ncu_sass_summary.py file.csv [kernel-substring]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2] if len(sys.argv) > 2 else None
# the file is a sequence of sections: ["Kernel Name", name], header row, data rows...
sections, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "data": []}
        sections.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = r
    elif cur is not None and len(r) == len(cur["hdr"]):
        cur["data"].append(r)
for sec in sections[:1] if want is None else [s for s in sections if want in s["name"]][:1]:
    hdr, data = sec["hdr"], sec["data"]
    iS, iE, iSm, iT = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
    tot = sum(int(r[iE]) for r in data)
    print(sec["name"], "static SASS:", len(data), "executed warp instr:", tot, "avg active threads: %.1f" % (sum(int(r[iT]) for r in data) / tot))
    ops, smp = collections.Counter(), collections.Counter()
    for r in data:
        toks = r[iS].split()
        op = toks[1] if toks[0].startswith("@") else toks[0]
        op = op.split(".")[0]
        ops[op] += int(r[iE])
        smp[op] += int(r[iSm] or 0)
    ts = sum(smp.values())
    print("%-10s %14s %7s %9s" % ("opcode", "executed", "share", "samples%"))
    for op, c in ops.most_common(28):
        print("%-10s %14d %6.1f%% %8.1f%%" % (op, c, 100 * c / tot, 100 * smp[op] / max(ts, 1)))
    print("async-copy / barrier opcodes (TMA evidence):")
    for r in data:
        if any(t in r[iS] for t in ("UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "UTCMMA", "TCGEN")) and int(r[iE]) > 0:
            print("  %10s  %s" % (r[iE], r[iS].strip()[:110]))
    print("stall samples:")
    for name in hdr:
        if name.startswith("stall_") and "Not Issued" not in name:
            i = hdr.index(name)
            s = sum(int(r[i] or 0) for r in data)
            if s:
                print("  %-24s %8d %5.1f%%" % (name, s, 100 * s / max(ts, 1)))
    print("hottest instructions by samples:")
    for r in sorted(data, key=lambda r: -int(r[iSm] or 0))[:25]:
        print("  %6s %10s  %s" % (r[iSm], r[iE], r[iS][:110]))
