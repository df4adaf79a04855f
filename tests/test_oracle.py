"""CPU: the oracle restatement (oracle/awi_oracle.c) against the committed outputs of the
reference itself (tests/golden, made by oracle/make_golden.py) and, where the reference
build is present, against the reference directly."""
import hashlib
import os

import numpy as np
import pytest

from conftest import golden_frames
from image_compression_b200 import synth


def sha(b):
    return hashlib.sha256(bytes(b)).hexdigest()


def test_golden_frames_encode_decode(port, golden):
    for i, kind, w, h, rows in golden_frames(golden):
        stride = synth.bmp_stride(w)
        payload = port.encode_bgr(rows, w, h, stride)
        assert payload.tobytes() == golden[f"payload_{i}"].tobytes(), (kind, w, h)
        bgr, used = port.decode(payload, w, h)
        assert used == payload.size
        assert bgr.tobytes() == golden[f"bgr_{i}"].tobytes(), (kind, w, h)
        rgb, _ = port.decode(payload, w, h, rgb_topdown=True)
        want = golden[f"bgr_{i}"].reshape(h, -1)[::-1, :3 * w].reshape(h, w, 3)[..., ::-1]
        assert rgb.tobytes() == want.tobytes()


def test_golden_hash_frames(port, golden):
    for line in golden["hashes"]:
        kind, w, h, seed, frame, n, hp, hb = str(line).split(",")
        w, h = int(w), int(h)
        rows = synth.frame_bgr_bottomup(kind, w, h, int(seed), int(frame))
        payload = port.encode_bgr(rows, w, h, synth.bmp_stride(w))
        assert payload.size == int(n) and sha(payload) == hp, (kind, w, h)
        bgr, _ = port.decode(payload, w, h)
        assert sha(bgr) == hb


def test_known_answer_blocks(port, golden):
    blocks = golden["kat_blocks"]
    off = 0
    for k, b in enumerate(blocks):
        assert (port.fdct(b) == golden["kat_coef"][k]).all()
        cl, cc = port.transform(b, 0), port.transform(b, 1)
        assert (cl == golden["kat_chunk_luma"][k]).all() and (cc == golden["kat_chunk_chroma"][k]).all()
        assert (port.inverse_transform(cl, 0) == golden["kat_inv_luma"][k]).all()
        assert (port.inverse_transform(cc, 1) == golden["kat_inv_chroma"][k]).all()
        n = int(golden["kat_rle_len"][k])
        want = golden["kat_rle_bytes"][off:off + n].tobytes()
        off += n
        prev = int(golden["kat_rle_prev"][k])
        assert port.rle(cl, prev) == want
        back, used = port.unrle(want, prev)
        assert used == n and (back == cl).all()


def test_survey_known_answers(port):
    """SURVEY.md A.7 (values observed from the reference's own functions)."""
    flat = {0: -1023, 1: -1015, 16: -895, 127: -7, 128: 0, 129: 7, 144: 127, 200: 575, 255: 1015}
    for v, dc in flat.items():
        c = port.fdct(np.full(64, v, np.uint8))
        assert c[0] == dc and not c[1:].any()
    yy, xx = np.mgrid[0:8, 0:8]
    ramp = (32 * yy + 4 * xx).astype(np.uint8)
    c = port.fdct(ramp).reshape(8, 8)
    assert list(c[0]) == [-15, -72, 0, -7, 0, -2, 0, 0]
    assert list(c[:, 0]) == [-15, -583, 0, -60, 0, -18, 0, -4]
    chunk = port.transform(ramp, 0)
    want = np.zeros(64, int)
    want[[0, 1, 17, 27, 29]] = [-1, -7, -49, -4, -1]
    assert (chunk == want).all()
    assert port.rle(chunk, 0).hex() == "ffe031ec749b110000"
    assert port.rle(chunk, int(chunk[0]) + 5).hex()[:6] == "ffa031"
    # quantiser ties, luma
    for c0, pos, q in ((8, 0, 1), (-8, 0, -1), (5, 2, 1), (6, 8, 1), (-6, 1, -1), (-18, 9, -2)):
        coef = np.zeros(64, np.int32)
        coef[pos] = c0
        assert port.quant(coef, 0)[pos] == q


def test_colour_samples_and_cube(port, golden):
    assert (port.rgb2ycc(golden["colour_in"]) == golden["colour_fwd"]).all()
    assert (port.ycc2rgb(golden["colour_in"]) == golden["colour_inv"]).all()
    a = np.arange(256, dtype=np.uint8)
    full = np.stack(np.meshgrid(a, a, a, indexing="ij"), -1).reshape(-1, 3)
    assert sha(port.rgb2ycc(full)) == str(golden["colour_cube_sha"][0])
    assert sha(port.ycc2rgb(full)) == str(golden["colour_cube_sha"][1])


def test_cos_table_hex(port):
    """SURVEY.md A.2."""
    t = port.cos_table().view(np.uint32)
    assert [hex(v) for v in t[0]] == ["0x3f800000", "0x3f7b14be", "0x3f6c835e", "0x3f54db31", "0x3f3504f3",
                                      "0x3f0e39da", "0x3ec3ef15", "0x3e47c5c2"]
    assert [hex(v) for v in t[1]] == ["0x3f800000", "0x3f54db31", "0x3ec3ef15", "0xbe47c5c2", "0xbf3504f3",
                                      "0xbf7b14be", "0xbf6c835e", "0xbf0e39da"]


def test_stripe_encode_matches_full(port):
    """awi_oracle_encode_rows: stripes with recomputed predictors concatenate to the full payload."""
    w, h = 200, 150
    rows = synth.frame_bgr_bottomup("photo", w, h, 0xA11CE05)
    stride = synth.bmp_stride(w)
    full = port.encode_bgr(rows, w, h, stride).tobytes()
    mbh = (h + 15) // 16
    for n in (2, 3, mbh):
        parts = [port.encode_rows(rows, w, h, stride, mbh * g // n, mbh * (g + 1) // n).tobytes() for g in range(n)]
        assert b"".join(parts) == full


def test_reference_sample1(port, ref, golden):
    """config[0]: sample/sample1.bmp golden round trip, digests recorded from the reference run."""
    path = "/root/reference/sample/sample1.bmp"
    local = os.path.join(os.path.dirname(__file__), "golden", "_local", "sample1.bmp")
    path = path if os.path.exists(path) else local
    if not os.path.exists(path):
        pytest.skip("reference sample image not on this machine")
    import struct
    bmp = open(path, "rb").read()
    n, h_awi, h_bmp, h_in = [str(x) for x in golden["sample1"]]
    assert sha(bmp) == h_in
    off = struct.unpack_from("<I", bmp, 10)[0]
    w, h = struct.unpack_from("<ii", bmp, 18)
    stride = synth.bmp_stride(w)
    rows = np.frombuffer(bmp, np.uint8)[off:off + stride * h]
    payload = port.encode_bgr(rows, w, h, stride)
    import oracle
    awi = oracle.awi_header(w, h) + payload.tobytes()
    assert len(awi) == int(n) and sha(awi) == h_awi
    assert awi[:48] == golden["sample1_prefix"].tobytes()
    bgr, _ = port.decode(payload, w, h)
    hdr = oracle.bmp_wrap(np.zeros(stride * h, np.uint8), w, h)[:54]
    assert sha(hdr + bgr.tobytes()) == h_bmp
    if ref is not None:
        assert ref.encode_bmp(bmp) == awi


def test_port_vs_reference_direct(port, ref):
    if ref is None:
        pytest.skip("reference build not present on this machine")
    rng = np.random.default_rng(3)
    for kind, w, h in (("noise", 33, 47), ("photo", 96, 80), ("flat", 16, 16), ("noise", 1, 1), ("photo", 16, 1)):
        rows = synth.frame_bgr_bottomup(kind, w, h, int(rng.integers(1 << 31)))
        a = ref.encode_bgr(rows, w, h)
        assert a.tobytes() == port.encode_bgr(rows, w, h, synth.bmp_stride(w)).tobytes()
        import oracle
        awi = oracle.awi_header(w, h) + a.tobytes()
        # row padding bytes are indeterminate in the reference (un-memset malloc, bmp.c:154)
        want = np.frombuffer(ref.decode_to_bmp(awi), np.uint8)[54:].reshape(h, -1)[:, :3 * w]
        assert (want == port.decode(a, w, h)[0].reshape(h, -1)[:, :3 * w]).all()
    for _ in range(300):
        b = rng.integers(0, 256, 64).astype(np.uint8)
        for ch in (0, 1):
            c = ref.transform(b, ch)
            assert (c == port.transform(b, ch)).all()
            assert (ref.inverse_transform(c, ch) == port.inverse_transform(c, ch)).all()


def test_chunk_starts_follow_eight_zero_bits(port):
    """The decoder's boundary discovery only parses from positions whose two preceding bytes hold 8 consecutive
    zero bits ending inside the last byte (the end-of-block symbol, rle.c:106, then at most 7 padding bits,
    rle.c:107).  Every real chunk start must satisfy that, whatever the content."""
    for kind, w, h, seed in (("photo", 320, 176, 3), ("noise", 160, 96, 4), ("flat", 64, 48, 5), ("photo", 37, 21, 6)):
        rows = synth.frame_bgr_bottomup(kind, w, h, synth.SEED_BASE + seed)
        p, _, lens = port.encode_bgr(rows, w, h, synth.bmp_stride(w), want_coefs=True)
        b = np.asarray(p, dtype=np.uint32)
        starts = np.cumsum(lens)[:-1].astype(np.int64)                  # every chunk start but the first
        assert (starts >= 3).all()
        win = (b[starts - 2] << 8) | b[starts - 1]
        z = ~win & 0xFFFF
        r = z & (z >> 1)
        r &= r >> 2
        r &= r >> 4
        assert ((r & 0xFF) != 0).all(), kind
        # and the predicate does prune: most positions fail it
        allpos = np.arange(3, b.size)
        win = (b[allpos - 2] << 8) | b[allpos - 1]
        z = ~win & 0xFFFF
        r = z & (z >> 1)
        r &= r >> 2
        r &= r >> 4
        if kind != "flat":
            assert ((r & 0xFF) != 0).mean() < 0.6


def test_band_helpers_equal_whole_frame(port, ref):
    """tests/_bands.py (the threaded whole-frame oracle runs the -m gpu tests use at 16384x16384 and
    8192x8192) gives the same bytes, lengths and pixels as the single-threaded whole-frame calls — on ragged
    sizes, with bands that do not divide the frame."""
    import _bands
    import oracle
    from image_compression_b200 import synth
    for kind, w, h in (("photo", 1000, 700), ("noise", 640, 427), ("photo", 2048, 1031)):
        stride = synth.bmp_stride(w)
        rows = synth.frame_bgr_bottomup(kind, w, h, 123)
        want, _, wlen = port.encode_bgr(rows, w, h, stride, want_coefs=True)
        p, lens = _bands.port_encode_frame(port, rows, w, h, stride, rows_per_band=5)
        assert p.tobytes() == want.tobytes() and (lens == wlen).all()
        offs = np.concatenate([[0], np.cumsum(lens.astype(np.int64))])[:-1]
        if ref is not None:
            _bands.ref_check_frame(ref, oracle, rows, w, h, stride, p, offs, rows_per_band=5)
            bad = p.copy()
            bad[offs[len(offs) // 2] + 2] ^= 0x10
            with pytest.raises(AssertionError):
                _bands.ref_check_frame(ref, oracle, rows, w, h, stride, bad, offs, rows_per_band=5)
        for rgb in (False, True):
            px = _bands.port_decode_frame(port, p, lens, w, h, stride, rgb, rows_per_band=5)
            assert px.tobytes() == port.decode(want, w, h, stride, rgb_topdown=rgb)[0].tobytes()
