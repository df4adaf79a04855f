#!/usr/bin/env python3
"""Generates image_compression_b200/csrc/codec_tables.h — every numeric constant the
kernels use, so none is typed by hand.

* COS_BITS: the reference's fp32 cosine table, (float)cos((2i+1)*j*PI/16) (dct.c:14-23),
  as bit patterns (pinned against SURVEY.md A.2 and the oracle in tests/).
* Quantiser tables (dct.c:86-107) and the scan gather table (dct.c:147-156).
* Fast-path constants for the AAN-scaled forward DCT:  for natural position k=(u,v)
    g = x*GK[k] - sign(x)*GC[k]   with x the AAN output,
    q = rn(g),  ambiguous  <=>  |g - q| > GT[k] = 0.5 - delta_k
  where GK = 1/(8 a_u a_v Q), a_0=1, a_j=sqrt2*cos(j*pi/16); GC = 1/(2Q) for odd Q, else 0;
  delta_k = 1.25*(E_k + R_k)/Q + 4e-5 with E_k the first-order fp32 rounding-error bound
  of the AAN network for |input| <= 128 (absolute-value network analysis below) and
  R_k the bound on |reference - ideal DCT| from the reference's own fp32 products
  and fp32 table rounding (DESIGN.md "exactness").
"""
import math
import os
import struct

import numpy as np

u = 2.0 ** -24

QL = [16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57, 69, 56,
      14, 17, 22, 29, 51, 87, 80, 62, 18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64, 81, 104, 113, 92,
      49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99]
QC = [17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99, 99, 99,
      47, 66, 99, 99, 99, 99, 99, 99] + [99] * 32
SCAN = [0, 1, 5, 6, 14, 15, 27, 28, 2, 4, 7, 13, 16, 26, 29, 42, 3, 8, 12, 17, 25, 30, 41, 43,
        9, 11, 18, 24, 31, 40, 44, 53, 10, 19, 23, 32, 39, 45, 52, 54, 20, 22, 33, 38, 46, 51, 55, 60,
        21, 34, 37, 47, 50, 56, 59, 61, 35, 36, 48, 49, 57, 58, 62, 63]


class Node:
    def __init__(self, M, E=0.0, integer=True):
        self.M, self.E, self.i = M, E, integer


def add(a, b):
    M = a.M + b.M
    if a.i and b.i and M <= 2 ** 24:
        return Node(M, 0.0, True)
    return Node(M, a.E + b.E + u * M, False)


def mulc(c, a):
    c = abs(c)
    M = c * a.M
    return Node(M, c * a.E + 2 * u * M, False)


def fmac(c, a, b):
    c = abs(c)
    M = c * a.M + b.M
    return Node(M, c * a.E + b.E + u * M + u * c * a.M, False)


def aan_abs(d):
    t0 = add(d[0], d[7]); t7 = add(d[0], d[7]); t1 = add(d[1], d[6]); t6 = add(d[1], d[6])
    t2 = add(d[2], d[5]); t5 = add(d[2], d[5]); t3 = add(d[3], d[4]); t4 = add(d[3], d[4])
    t10 = add(t0, t3); t13 = add(t0, t3); t11 = add(t1, t2); t12 = add(t1, t2)
    o = [None] * 8
    o[0] = add(t10, t11); o[4] = add(t10, t11)
    z1 = mulc(0.707106781, add(t12, t13))
    o[2] = add(t13, z1); o[6] = add(t13, z1)
    t10 = add(t4, t5); t11 = add(t5, t6); t12 = add(t6, t7)
    z5 = mulc(0.382683433, add(t10, t12))
    z2 = fmac(0.541196100, t10, z5)
    z4 = fmac(1.306562965, t12, z5)
    z3 = mulc(0.707106781, t11)
    z11 = add(t7, z3); z13 = add(t7, z3)
    o[5] = add(z13, z2); o[3] = add(z13, z2); o[1] = add(z11, z4); o[7] = add(z11, z4)
    return o


def error_bounds():
    rows = [aan_abs([Node(128.0) for _ in range(8)]) for _ in range(8)]
    a = [1.0] + [math.sqrt(2) * math.cos(k * math.pi / 16) for k in range(1, 8)]
    sabs = [sum(abs(math.cos((2 * i + 1) * k * math.pi / 16)) for i in range(8)) for k in range(8)]
    E = np.zeros(64)
    for v in range(8):
        col = aan_abs([rows[y][v] for y in range(8)])
        for uu in range(8):
            K = 1 / (8 * a[uu] * a[v])
            sc = 0.25 * (2 ** -0.5 if uu == 0 else 1) * (2 ** -0.5 if v == 0 else 1)
            R = sc * 128 * 4 * u * sabs[uu] * sabs[v]
            E[uu * 8 + v] = K * col[uu].E + R
    return E, a


def colour_fix_tables():
    """Where the reference's fp64 colour sums land below an integer that the exact rational reaches.

    color.c:31-33 evaluates Y = 0.299 r + 0.587 g + 0.114 b (and Cb, Cr likewise, + 128) left to right in fp64 and
    truncates.  The rational value is n / 1000 (n / 10000) with an integer numerator n, so floor(n / 1000) is the
    reference's result unless n is an exact multiple: then fp64 may end one ulp below the integer and truncation gives
    one less.  For a given (r, g) at most one b makes the Y (or Cr) numerator a multiple, so one bit per (r, g) says
    whether that happens: 65536 bits per table, bit index (r << 8) | g.  Cb never differs (checked here).
    numpy's float64 arithmetic is IEEE, one rounding per operation, no contraction: the reference's -O0 build."""
    r, g, b = np.meshgrid(np.arange(256, dtype=np.int64), np.arange(256, dtype=np.int64), np.arange(256, dtype=np.int64), indexing="ij")
    rf, gf, bf = r.astype(np.float64), g.astype(np.float64), b.astype(np.float64)

    def clamp_trunc(v):
        return np.clip(np.trunc(v), 0, 255).astype(np.int64)

    tabs = {}
    specs = {
        "Y": (299 * r + 587 * g + 114 * b, 1000, clamp_trunc((0.299 * rf + 0.587 * gf) + 0.114 * bf)),
        "CB": (-1687 * r - 3313 * g + 5000 * b + 1280000, 10000, clamp_trunc(((-0.1687 * rf - 0.3313 * gf) + 0.5 * bf) + 128.0)),
        "CR": (5000 * r - 4187 * g - 813 * b + 1280000, 10000, clamp_trunc(((0.5 * rf - 0.4187 * gf) - 0.0813 * bf) + 128.0)),
    }
    for name, (n, den, ref) in specs.items():
        q = n // den
        diff = q - ref
        assert ((diff == 0) | ((diff == 1) & (n % den == 0))).all(), name      # only exact multiples differ, and by one
        bits = (diff == 1)
        assert (bits.sum(axis=2) <= 1).all()
        if name == "CB":
            assert not bits.any()                                           # Cb needs no table
            continue
        flat = bits.any(axis=2).reshape(-1)                                  # index (r << 8) | g
        words = np.zeros(2048, np.uint32)
        idx = np.nonzero(flat)[0]
        np.bitwise_or.at(words, idx >> 5, (np.uint32(1) << (idx & 31).astype(np.uint32)))
        tabs[name] = (words, int(bits.sum()))
    return tabs


def write_colour_fix(path):
    tabs = colour_fix_tables()
    out = ["// GENERATED by tools/gen_tables.py — do not edit.",
           "// Exact-multiple corrections of the forward colour transform (color.c:31-33): bit (r << 8 | g) of the table is set",
           "// when the one b that makes the numerator an exact multiple of the denominator gives a reference result one below",
           "// the rational quotient (fp64 lands just under the integer).  Cb has no such case.",
           "#pragma once", "#include <stdint.h>"]
    for name, (words, cnt) in tabs.items():
        out.append("// %d of 2^24 colours" % cnt)
        out.append("#define AWI_%s_FIX_BITS { \\" % name)
        for i in range(0, 2048, 16):
            out.append("    " + ", ".join("0x%08xu" % w for w in words[i:i + 16]) + (", \\" if i + 16 < 2048 else " }"))
    with open(path, "w") as f:
        f.write("\n".join(out) + "\n")
    print("wrote", os.path.normpath(path))


# ---- the same analysis with nothing dropped (the bound the exactness argument rests on).
# A node carries M >= |ideal value| and E >= |computed - ideal|.  fl() is fp32 round-to-nearest: |fl(x) - x| <= u |x|.
# Constants are fp32 roundings of 9-digit decimals of the exact factors: relative error <= UC.
UC = 1.01 * u


def r_add(a, b):
    M = a.M + b.M
    if a.i and b.i and M <= 2 ** 24:
        return Node(M, 0.0, True)                      # integers below 2^24: the fp32 sum is exact
    return Node(M, (a.E + b.E) * (1 + u) + u * M, False)        # |computed operands' sum| <= M + Ea + Eb


def r_mulc(c, a):
    c = abs(c)
    # fl(c^ a^), c^ = c (1 + d), |d| <= UC:  |c^ a^ - c a| <= c (1 + UC) Ea + UC c Ma, then one rounding of at most that much more than c Ma
    inner = c * (1 + UC) * a.E + UC * c * a.M
    return Node(c * a.M, inner * (1 + u) + u * c * a.M, False)


def r_fmac(c, a, b):
    c = abs(c)
    inner = c * (1 + UC) * a.E + UC * c * a.M + b.E
    return Node(c * a.M + b.M, inner * (1 + u) + u * (c * a.M + b.M), False)


def aan_rig(d):
    add, mulc, fmac = r_add, r_mulc, r_fmac
    t0 = add(d[0], d[7]); t7 = add(d[0], d[7]); t1 = add(d[1], d[6]); t6 = add(d[1], d[6])
    t2 = add(d[2], d[5]); t5 = add(d[2], d[5]); t3 = add(d[3], d[4]); t4 = add(d[3], d[4])
    t10 = add(t0, t3); t13 = add(t0, t3); t11 = add(t1, t2); t12 = add(t1, t2)
    o = [None] * 8
    o[0] = add(t10, t11); o[4] = add(t10, t11)
    z1 = mulc(0.707106781, add(t12, t13))
    o[2] = add(t13, z1); o[6] = add(t13, z1)
    t10 = add(t4, t5); t11 = add(t5, t6); t12 = add(t6, t7)
    z5 = mulc(0.382683433, add(t10, t12))
    z2 = fmac(0.541196100, t10, z5)
    z4 = fmac(1.306562965, t12, z5)
    z3 = mulc(0.707106781, t11)
    z11 = add(t7, z3); z13 = add(t7, z3)
    o[5] = add(z13, z2); o[3] = add(z13, z2); o[1] = add(z11, z4); o[7] = add(z11, z4)
    return o


def error_bounds_rigorous():
    """|fast path - reference| before truncation, per position, for every block of samples 0..255: the AAN network's
    accumulated fp32 error (scaled to the reference's units) plus the reference's own distance from the ideal DCT
    (two fp32 roundings per product and the fp32 cosine table: (1 + u)^4 - 1 per term; its fp64 sum is exact and its
    fp64 scaling adds 2^-52 relative)."""
    rows = [aan_rig([Node(128.0) for _ in range(8)]) for _ in range(8)]
    a = [1.0] + [math.sqrt(2) * math.cos(k * math.pi / 16) for k in range(1, 8)]
    sabs = [sum(abs(math.cos((2 * i + 1) * k * math.pi / 16)) for i in range(8)) for k in range(8)]
    E = np.zeros(64)
    for v in range(8):
        col = aan_rig([rows[y][v] for y in range(8)])
        for uu in range(8):
            K = 1 / (8 * a[uu] * a[v])
            sc = 0.25 * (2 ** -0.5 if uu == 0 else 1) * (2 ** -0.5 if v == 0 else 1)
            R = sc * 128 * ((1 + u) ** 4 - 1 + 2.0 ** -50) * sabs[uu] * sabs[v]
            E[uu * 8 + v] = K * col[uu].E + R
    return E


# The quantiser's own arithmetic (codec_math.cuh quant_absdiff): g^ = fl(x gk^ + sc), gk^ and |sc| = gc^ fp32 roundings
# of K/Q and 1/(2Q), |g| <= 128:  |g^ - (x K / Q - sign(x)/(2Q))| <= (2 |g| + 1/(2Q)) u (1 + u) + |g| u  <  QUANT_ARITH.
QUANT_ARITH = 3.2 * 128 * u


def f32(x):
    return float(np.float32(x))


def flit(x):
    return "%.9ef" % f32(x)


def main():
    E, a = error_bounds()
    E_rig = error_bounds_rigorous()
    # What the thresholds need: delta_k >= (rigorous bound) / Q + the quantiser's own error.
    # delta_k below is 1.25 x the first-order bound / Q + 4e-5; the margin is checked here, for both tables.
    for Qt in (QL, QC):
        for k in range(1, 64):
            need = E_rig[k] / Qt[k] + QUANT_ARITH
            have = 1.25 * E[k] / Qt[k] + 4e-5
            assert need <= have, (k, need, have)
    assert (E_rig >= E).all() and (E_rig <= 1.01 * E).all()
    out = []
    out.append("// GENERATED by tools/gen_tables.py — do not edit.")
    out.append("#pragma once")
    out.append("#include <stdint.h>")
    cosb = []
    for i in range(8):
        for j in range(8):
            val = np.float32(math.cos(float((2 * i + 1) * j) * 3.14159265358979323846 / 16.0))
            cosb.append(struct.unpack("<I", struct.pack("<f", val))[0])
    out.append("// (float)cos((2i+1)*j*PI/16), [spatial i][frequency j]  (dct.c:14-23)")
    out.append("#define AWI_COS_BITS { " + ", ".join("0x%08xu" % b for b in cosb) + " }")
    out.append("#define AWI_QLUMA { " + ", ".join(map(str, QL)) + " }")
    out.append("#define AWI_QCHROMA { " + ", ".join(map(str, QC)) + " }")
    out.append("#define AWI_SCAN { " + ", ".join(map(str, SCAN)) + " }")
    inv = [0] * 64
    for i, s in enumerate(SCAN):
        inv[s] = i
    out.append("// inverse of AWI_SCAN: natural position -> index in the chunk")
    out.append("#define AWI_UNSCAN { " + ", ".join(map(str, inv)) + " }")
    for name, Q in (("LUMA", QL), ("CHROMA", QC)):
        gk, gc, gt, dl = [], [], [], []
        for k in range(64):
            uu, v = divmod(k, 8)
            K = 1 / (8 * a[uu] * a[v])
            gk.append(flit(K / Q[k]))
            gc.append(flit(1 / (2 * Q[k]) if Q[k] % 2 else 0.0))
            delta = 1.25 * E[k] / Q[k] + 4e-5
            dl.append(delta)
            gt.append(flit(0.5 - delta))
        out.append("#define AWI_GK_%s { %s }" % (name, ", ".join(gk)))
        out.append("#define AWI_GC_%s { %s }" % (name, ", ".join(gc)))
        out.append("#define AWI_GT_%s { %s }" % (name, ", ".join(gt)))
        out.append("// max delta %s = %.3e" % (name, max(dl)))
        out.append("// one threshold for every position (the tightest): costs a few more exact recomputations, saves 8 registers")
        out.append("#define AWI_GT_MIN_%s %s" % (name, flit(0.5 - max(dl))))
    out.append("// T-space error bound per position (fast path vs reference), max %.4e; with every second-order term kept the"
               " bound is at most %.6f x this (tools/gen_tables.py error_bounds_rigorous), well inside the 1.25 x the thresholds use" % (E.max(), float((E_rig / E).max())))
    out.append("#define AWI_EPS_T { " + ", ".join("%.6e" % e for e in E) + " }")
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "image_compression_b200", "csrc",
                        "codec_tables.h")
    with open(path, "w") as f:
        f.write("\n".join(out) + "\n")
    print("wrote", os.path.normpath(path))
    write_colour_fix(os.path.join(os.path.dirname(path), "colour_fix_tables.h"))


if __name__ == "__main__":
    main()
