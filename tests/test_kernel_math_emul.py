"""CPU: the arithmetic the CUDA kernels run (csrc/codec_math.cuh compiled for the host by
tests/emul) against the oracle.  This is where the exactness scheme is proven without a GPU:
the fast AAN path may only differ from the reference where it raised an ambiguity flag."""
import ctypes as C

import numpy as np

u8p = C.POINTER(C.c_uint8)
i32p = C.POINTER(C.c_int)


def _blocks_adversarial():
    yy, xx = np.mgrid[0:8, 0:8]
    out = []
    for u in range(8):
        for v in range(8):
            pat = np.sign(np.cos((2 * yy + 1) * u * np.pi / 16) * np.cos((2 * xx + 1) * v * np.pi / 16))
            for lo, hi in ((0, 255), (1, 254), (127, 129), (0, 128), (128, 255), (100, 156)):
                out.append(np.where(pat > 0, hi, lo).ravel())
                out.append(np.where(pat > 0, lo, hi).ravel())
    for v in range(256):
        out.append(np.full(64, v))
    return np.array(out, np.uint8)


def _transform(emul, blocks, chroma):
    blocks = np.ascontiguousarray(blocks, np.uint8)
    out = np.empty((len(blocks), 64), np.int32)
    n = emul.emul_transform_n(blocks.ctypes.data_as(u8p), C.c_long(len(blocks)), chroma, out.ctypes.data_as(i32p))
    return out, n


def test_transform_bit_exact(emul, port):
    rng = np.random.default_rng(11)
    yy, xx = np.mgrid[0:8, 0:8]
    smooth = []
    for _ in range(4000):
        a, b = rng.uniform(-12, 12, 2)
        smooth.append(np.clip(rng.uniform(0, 255) + a * xx + b * yy + rng.integers(-3, 4, (8, 8)), 0, 255).ravel())
    sets = {"noise": rng.integers(0, 256, (4000, 64)), "smooth": np.array(smooth), "adversarial": _blocks_adversarial(),
            "near-flat": np.clip(rng.normal(128, 2, (2000, 64)), 0, 255)}
    for name, blocks in sets.items():
        blocks = blocks.astype(np.uint8)
        for chroma in (0, 1):
            got, n_exact = _transform(emul, blocks, chroma)
            want = np.stack([port.transform(b, chroma) for b in blocks])
            assert (got == want).all(), (name, chroma)
            # the fallback must stay rare or the fast path is pointless
            assert n_exact <= 0.002 * blocks.size, (name, chroma, n_exact)


def test_fast_path_only_differs_where_flagged(emul, port):
    rng = np.random.default_rng(12)
    blocks = rng.integers(0, 256, (3000, 64)).astype(np.uint8)
    from oracle import port as _p
    _, _, scan = _p().tables()
    for chroma in (0, 1):
        for b in blocks:
            chunk = np.empty(64, np.int32)
            amb = C.c_uint64()
            emul.emul_fast_only(b.ctypes.data_as(u8p), chroma, chunk.ctypes.data_as(i32p), C.byref(amb))
            want = port.transform(b, chroma)
            bad = np.nonzero(chunk != want)[0]
            for i in bad:
                assert (amb.value >> int(scan[i])) & 1, "unflagged fast-path mismatch"
                assert abs(int(chunk[i]) - int(want[i])) == 1


def test_dc_formula_exhaustive(emul, port):
    """Every block sum S in [-8192, 8128]: integer DC == dct_block's DC, both tables."""
    ql, qc, _ = port.tables()
    for S in range(-8192, 8129):
        # a block with that sum: fill with floor, distribute the remainder
        base, rem = divmod(S + 128 * 64, 64)
        blk = np.full(64, base, np.int64)
        blk[:rem] += 1
        if blk.min() < 0 or blk.max() > 255:
            continue
        if S % 97 and abs(S) > 64:      # full pipeline on a thinned set, the closed form on all
            continue
        c = port.fdct(blk.astype(np.uint8))[0]
        for chroma, Q in ((0, ql[0]), (1, qc[0])):
            coef = np.zeros(64, np.int32)
            coef[0] = c
            assert emul.emul_dc_from_sum(S, chroma) == port.quant(coef, chroma)[0], S


def test_quant_int_matches_reference_rounding(emul, port):
    ql, qc, _ = port.tables()
    for chroma, tab in ((0, ql), (1, qc)):
        for c in range(-1100, 1101):
            want = port.quant(np.full(64, c, np.int32), chroma)
            got = np.array([emul.emul_quant_int(c, int(q)) for q in tab])
            assert (got == want).all(), c


def test_rle_matches(emul, port, golden):
    rng = np.random.default_rng(13)
    chunks = [c.astype(np.int32) for c in golden["kat_chunk_luma"]]
    for _ in range(500):   # sparse, dense, long zero runs, extreme magnitudes
        c = np.zeros(64, np.int32)
        k = rng.integers(0, 64)
        idx = rng.choice(63, k, replace=False) + 1
        c[idx] = rng.integers(-95, 96, k)
        c[0] = rng.integers(-64, 65)
        chunks.append(c)
    c = np.zeros(64, np.int32); c[63] = 1; chunks.append(c.copy())           # 3 ZRLs then a value
    c = np.zeros(64, np.int32); c[17] = -1; c[34] = 1; c[51] = -95; chunks.append(c.copy())
    c = np.full(64, 95, np.int32); c[0] = 64; chunks.append(c.copy())           # longest chunk: 121 bytes
    c = np.full(64, -95, np.int32); c[0] = -64; chunks.append(c.copy())
    out = np.empty(256, np.uint8)
    for c in chunks:
        for prev in (0, int(rng.integers(-64, 65)), 64, -64):
            c = np.ascontiguousarray(c, np.int32)
            n = emul.emul_rle(c.ctypes.data_as(i32p), prev, out.ctypes.data_as(u8p))
            assert out[:n].tobytes() == port.rle(c, prev), (c, prev)
    assert len(port.rle(np.array([64] + [95] * 63, np.int32), -64)) == 121


def test_colour_exhaustive(emul, port):
    """All 2^24 inputs, forward and inverse."""
    a = np.arange(256, dtype=np.uint8)
    cube = np.ascontiguousarray(np.stack(np.meshgrid(a, a, a, indexing="ij"), -1).reshape(-1, 3))
    out = np.empty_like(cube)
    emul.emul_rgb2ycc_n(cube.ctypes.data_as(u8p), C.c_long(len(cube)), out.ctypes.data_as(u8p))
    assert (out == port.rgb2ycc(cube)).all()
    emul.emul_ycc2rgb_n(cube.ctypes.data_as(u8p), C.c_long(len(cube)), out.ctypes.data_as(u8p))
    assert (out == port.ycc2rgb(cube)).all()


def test_generated_tables_match_oracle(emul, port):
    bits = np.empty(64, np.uint32)
    emul.emul_cos_bits(bits.ctypes.data_as(C.POINTER(C.c_uint32)))
    assert (bits == port.cos_table().view(np.uint32).ravel()).all()


def test_colour_fix_tables_reproduce_the_reference_on_every_colour(port):
    """csrc/colour_fix_tables.h (generated): integer quotient minus the table bit == the reference's fp64 result
    (color.c:31-33 + utils.c:10-17) for all 2^24 (r, g, b), which is what the encode kernel's colour phase computes;
    Cb is the plain quotient."""
    import os
    import re
    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    path = os.path.join(ROOT, "image_compression_b200", "csrc", "colour_fix_tables.h")
    text = open(path).read().replace("\\\n", " ")
    tabs = {}
    for name in ("Y", "CR"):
        body = re.search(r"#define AWI_%s_FIX_BITS \{(.*?)\}" % name, text, re.S).group(1)
        tabs[name] = np.array([int(x, 16) for x in re.findall(r"0x([0-9a-f]{8})u", body)], np.uint32)
        assert tabs[name].size == 2048
    for r0 in range(0, 256, 32):                       # 2^21 colours at a time
        r, g, b = np.meshgrid(np.arange(r0, r0 + 32, dtype=np.int64), np.arange(256, dtype=np.int64), np.arange(256, dtype=np.int64), indexing="ij")
        rgb = np.stack([r, g, b], axis=-1).reshape(-1, 3).astype(np.uint8)
        want = port.rgb2ycc(rgb).astype(np.int64)
        r, g, b = (rgb[:, i].astype(np.int64) for i in range(3))
        idx = (r << 8) | g

        def bit(t):
            return (t[idx >> 5].astype(np.int64) >> (idx & 31)) & 1

        ny = 299 * r + 587 * g + 114 * b
        y = ny // 1000 - np.where(ny % 1000 == 0, bit(tabs["Y"]), 0)
        ncb = -1687 * r - 3313 * g + 5000 * b + 1280000
        ncr = 5000 * r - 4187 * g - 813 * b + 1280000
        cr = ncr // 10000 - np.where(ncr % 10000 == 0, bit(tabs["CR"]), 0)
        assert (y == want[:, 0]).all() and (ncb // 10000 == want[:, 1]).all() and (cr == want[:, 2]).all()


def _eps_t():
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "image_compression_b200", "csrc", "codec_tables.h")).read()
    body = re.search(r"#define AWI_EPS_T \{(.*?)\}", text).group(1)
    return np.array([float(x) for x in body.split(",")])


def test_fast_path_stays_inside_the_a_priori_bound(emul):
    """|fast AAN value - reference's exact pre-truncation value| <= AWI_EPS_T[k] at every position, on the blocks that
    push each coefficient to its largest magnitude (sign patterns at full swing), on flat blocks and on noise.  The bound
    is an a-priori worst case (tools/gen_tables.py: every rounding of the butterfly network at its largest possible
    operand, second-order terms kept, plus the reference's own distance from the ideal DCT); the ambiguity thresholds are
    built from 1.25 x it, so the measured maximum has to sit well below 1."""
    eps = _eps_t()
    a = np.array([1.0] + [np.sqrt(2) * np.cos(k * np.pi / 16) for k in range(1, 8)])
    K = 1.0 / (8.0 * np.outer(a, a)).ravel()
    rng = np.random.default_rng(21)
    blocks = np.concatenate([_blocks_adversarial(), rng.integers(0, 256, (3000, 64)).astype(np.uint8),
                             np.where(rng.integers(0, 2, (3000, 64)) > 0, 255, 0).astype(np.uint8)])
    f64p = C.POINTER(C.c_double)
    worst = np.zeros(64)
    for b in blocks:
        b = np.ascontiguousarray(b, np.uint8)
        fast, exact = np.empty(64), np.empty(64)
        emul.emul_fast_vs_exact(b.ctypes.data_as(u8p), fast.ctypes.data_as(f64p), exact.ctypes.data_as(f64p))
        worst = np.maximum(worst, np.abs(fast * K - exact))
    assert (worst <= eps).all(), (worst / eps).max()
    assert (worst / eps).max() < 0.8            # the bound is a worst case over all roundings at once: nothing gets close


def test_kernel_flag_expression_is_safe(emul, port):
    """The encode kernel tests a running maximum against ONE threshold (the tightest of the per-position table) and, only
    where that trips, every position against its own: the flags must come out as the per-position flags exactly, and what
    is not flagged must already equal the reference."""
    rng = np.random.default_rng(22)
    from oracle import port as _p
    _, _, scan = _p().tables()
    blocks = np.concatenate([_blocks_adversarial(), rng.integers(0, 256, (2500, 64)).astype(np.uint8)])
    for chroma in (0, 1):
        for b in blocks:
            b = np.ascontiguousarray(b, np.uint8)
            chunk, chunk2 = np.empty(64, np.int32), np.empty(64, np.int32)
            amb, amb2 = C.c_uint64(), C.c_uint64()
            emul.emul_fast_kernel_flags(b.ctypes.data_as(u8p), chroma, chunk.ctypes.data_as(i32p), C.byref(amb))
            emul.emul_fast_only(b.ctypes.data_as(u8p), chroma, chunk2.ctypes.data_as(i32p), C.byref(amb2))
            assert (chunk == chunk2).all() and amb2.value == amb.value
            want = port.transform(b, chroma)
            for i in np.nonzero(chunk != want)[0]:
                assert (amb.value >> int(scan[i])) & 1, "unflagged fast-path mismatch under the kernel's flag expression"


def test_invariant_divisor_division(emul):
    """fast_div (one multiply-high and a shift in the encode kernel's prologue) equals integer division for every
    divisor a frame can produce (macroblocks per row <= 4096, tiles per frame < 2^20) and numerators below 2^31."""
    rng = np.random.default_rng(5)
    ns = np.concatenate([np.arange(0, 70), rng.integers(0, 2 ** 31, 400), [2 ** 31 - 1, 2 ** 31 - 2, 2 ** 30, 2 ** 30 - 1]])
    ns = np.ascontiguousarray(ns, np.int32)
    bad = C.c_int32()
    emul.emul_fastdiv_check.restype = C.c_uint32
    nsp = ns.ctypes.data_as(i32p)
    for d0, d1 in ((1, 70000), (2 ** 20 - 500, 2 ** 20 + 500), (2 ** 30 - 50, 2 ** 30 + 1)):
        d = emul.emul_fastdiv_check(C.c_uint32(d0), C.c_uint32(d1), nsp, len(ns), C.byref(bad))
        assert d == 0, (d, bad.value)


def test_cosine_table_first_column_is_one(emul):
    """dec_reconstruct starts its sums from the DC product instead of adding it in a loop turn: valid because
    cos_table[.][0] (dct.c:14-23) is exactly 1.0f, so ((cu*cv*coef) * 1.0) * 1.0 is the product itself."""
    bits = np.zeros(64, np.uint32)
    emul.emul_cos_bits(bits.ctypes.data_as(C.POINTER(C.c_uint32)))
    assert (bits.reshape(8, 8)[:, 0] == 0x3F800000).all()
