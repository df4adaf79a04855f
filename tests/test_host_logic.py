"""CPU: host-side logic — synthetic generators, BMP/AWI container mirror, stripe partition
and stitching (including a 2-rank gloo run in which the oracle stands in for the device)."""
import io
import os
import struct
import subprocess
import sys

import numpy as np

from image_compression_b200 import awi, bmp, stripes, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_synth_layout_and_determinism():
    w, h = 37, 21
    rows = synth.frame_bgr_bottomup("photo", w, h, 0xA11CE01, 3)
    assert rows.shape == (h, synth.bmp_stride(w)) and not rows[:, 3 * w:].any()
    rgb = synth.frame_rgb_topdown("photo", w, h, 0xA11CE01, 3)
    assert (rows[h - 1 - 5, 3 * 7: 3 * 7 + 3] == rgb[5, 7, ::-1]).all()
    assert (synth.frame_rgb_topdown("noise", w, h, 1) != synth.frame_rgb_topdown("noise", w, h, 2)).any()
    assert (synth.frame_rgb_topdown("flat", 4, 4, 0) == 144).all()
    # noise byte = fmix32(i*0x9E3779B1 + seed) >> 24 at i = (y*w+x)*3+c
    i = (5 * w + 7) * 3 + 1
    hsh = (i * 0x9E3779B1 + 0xA11CE01) & 0xFFFFFFFF
    hsh ^= hsh >> 16; hsh = hsh * 0x85EBCA6B & 0xFFFFFFFF; hsh ^= hsh >> 13
    hsh = hsh * 0xC2B2AE35 & 0xFFFFFFFF; hsh ^= hsh >> 16
    assert synth.frame_rgb_topdown("noise", w, h, 0xA11CE01)[5, 7, 1] == hsh >> 24


def test_stripe_partition_and_windows():
    for mbh in (1, 7, 68, 1024):
        for n in (1, 2, 3, 4, 8):
            parts = stripes.stripe_rows(mbh, n)
            assert parts[0][0] == 0 and parts[-1][1] == mbh
            assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
            assert max(r1 - r0 for r0, r1 in parts) - min(r1 - r0 for r0, r1 in parts) <= 1
    assert stripes.stripe_window(1080, 0, 34) == (0, 544)
    assert stripes.stripe_window(1080, 34, 68) == (528, 1080)
    assert stripes.window_byte_range(1080, 5760, 528, 1080) == (0, 552 * 5760)
    assert stripes.segment_offsets([5, 0, 7]) == [0, 5, 5]
    assert list(stripes.batch_share(10, 1, 4)) == [1, 5, 9]


def test_bmp_reader_padding_and_offsets():
    """bmp.c:48-119 semantics of the Python mirror: bfOffBits honoured, bottom-up, zero padding."""
    w, h = 21, 19
    rows = synth.frame_bgr_bottomup("noise", w, h, 5)
    import oracle
    body = oracle.bmp_wrap(rows, w, h)
    # move the pixel array to offset 138 like the reference's sample (BITMAPV5 header)
    shifted = body[:10] + struct.pack("<I", 138) + body[14:54] + b"\0" * 84 + body[54:]
    pi = bmp.BMPPixelInput(io.BytesIO(shifted))
    assert pi.begin() and (pi.width, pi.height) == (w, h)
    got, stride = pi.rows_bgr_bottomup()
    assert stride == synth.bmp_stride(w) and got.tobytes() == rows.tobytes()
    blk = bmp.RGBBlock16()
    assert pi.read_block(1, 1, blk)
    rgb = synth.frame_rgb_topdown("noise", w, h, 5)
    assert (blk.r[:3, :5] == rgb[16:19, 16:21, 0]).all() and not blk.r[3:].any() and not blk.g[:, 5:].any()


def test_awi_header_and_error_paths(capsys):
    assert awi.pack_header(640, 427).hex() == "415745534f4d454980020000ab010000"

    class Sink(bmp.PixelOutput):
        pass
    assert awi.decompress_awi(awi.AWIFileInput(io.BytesIO(b"NOTANAWI" + b"\0" * 8)), Sink()) is False
    assert "Invalid file format!" in capsys.readouterr().err
    assert awi.decompress_awi(awi.AWIFileInput(io.BytesIO(b"AWE")), Sink()) is False
    assert "Reading header error!" in capsys.readouterr().err


def test_bmp_writer_layout():
    w, h = 5, 3
    out = io.BytesIO()
    po = bmp.BMPPixelOutput(out)
    assert po.begin(w, h)
    blk = bmp.RGBBlock16()
    blk.r[:h, :w], blk.g[:h, :w], blk.b[:h, :w] = 1, 2, 3
    blk.r[0, 0] = 9
    assert po.write_block(0, 0, blk) and po.end()
    data = out.getvalue()
    assert len(data) == 54 + 16 * 3 and data[:2] == b"BM"
    assert struct.unpack_from("<I", data, 10)[0] == 54 and struct.unpack_from("<ii", data, 18) == (w, h)
    top = data[54 + 2 * 16: 54 + 3 * 16]          # top image row is the LAST file row
    assert top[:3] == bytes([3, 2, 9]) and top[15] == 0


def test_two_rank_stripe_stitch_gloo(tmp_path):
    """world_size 2 over gloo: each rank encodes its macroblock-row stripe (oracle as the
    device stand-in), lengths are all-gathered, rank 0 stitches == full-frame payload."""
    script = os.path.join(ROOT, "tests", "_gloo_stripes.py")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29517")
    procs = [subprocess.Popen([sys.executable, script, str(r), "2", str(tmp_path)], env=env, stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=180)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "STITCH_OK" in outs[0]
