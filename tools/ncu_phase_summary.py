#!/usr/bin/env python3
"""Executed warp instructions / stall samples per kernel phase of awi_encode.cu (phases are
delimited by the '---- X:' comments and section markers).  Usage: ncu_phase_summary.py cuda_sass.csv"""
import collections
import csv
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rows = list(csv.reader(open(sys.argv[1])))
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
            out.append((fname, int(r[0]), int(r[hdr.index("Instructions Executed")]), int(r[hdr.index("# Samples")])))
        except ValueError:
            pass
enc = open(os.path.join(ROOT, "image_compression_b200/csrc/awi_encode.cu")).read().split("\n")
mth = open(os.path.join(ROOT, "image_compression_b200/csrc/codec_math.cuh")).read().split("\n")


def marks(lines, pats):
    res = []
    for name, pat in pats:
        for i, l in enumerate(lines):
            if pat in l:
                res.append((name, i + 1))
                break
    return res


M_ENC = marks(enc, [("top", "#include"), ("lookback helpers", "Look-back descriptors"), ("loader fn", "phase A"), ("colour helpers", "phase B helpers"),
                    ("luma_piece", "uint4 luma_piece("), ("chroma_piece", "uint4 chroma_piece("), ("halo colour", "one colour-difference sample (halo"),
                    ("sample addressing", "Sample bytes of a tile"), ("prologue", "the kernel"), ("A stage", "---- A:"), ("B1 colour loop", "---- B1:"),
                    ("B2 dct/quant loop", "---- B2:"), ("halo", "halo macroblock: only"), ("C exact", "---- C:"), ("D rle", "---- D:"),
                    ("E scan", "---- E:"), ("two-pass tail", "if (two_pass) {"), ("lookback", "---- single-pass tail"), ("F copy", "---- F:"),
                    ("gather kernels", "two-pass mode, passes 2 and 3"), ("launch", "-- launch")])
M_MTH = marks(mth, [("bit helpers", "bit helpers"), ("colour fwd", "forward colour"), ("colour inv", "inverse colour"), ("centred", "sample (0..255)"),
                    ("aan", "forward DCT, fast path"), ("dc/quant_int", "Quantised DC from"), ("quant", "Quantise one AAN-scaled"),
                    ("fdct_quant_pack", "Fast DCT + quantise"), ("exact path", "forward DCT, exact path"), ("rle", "RLE of one chunk")])


def bucket(f, l):
    m = M_ENC if f == "awi_encode.cu" else M_MTH if f == "codec_math.cuh" else None
    if m is None:
        return f
    name = m[0][0]
    for n, s in m:
        if l >= s:
            name = n
    return f.split(".")[0][:10] + ":" + name


E, S = collections.Counter(), collections.Counter()
for f, l, e, s in out:
    E[bucket(f, l)] += e
    S[bucket(f, l)] += s
te, ts = sum(E.values()), sum(S.values())
for k, v in E.most_common():
    print("%-36s exec %5.1f%%   samples %5.1f%%" % (k, 100 * v / te, 100 * S[k] / ts))
