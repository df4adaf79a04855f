"""GPU (-m gpu): the C host side (host/awi_gpu.c, host/awi_cli.c) — the reference's
compress_to_awi / decompress_awi signatures over the C ABI, driven the way the reference's
main.c drives them, checked against the oracle."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import oracle
from image_compression_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def host():
    from image_compression_b200 import build
    build.build()
    subprocess.run(["make", "-C", os.path.join(ROOT, "host")], check=True, stdout=subprocess.DEVNULL)
    return os.path.join(ROOT, "host")


def test_cli_convert_revert_show(host, port, tmp_path):
    w, h = 333, 201
    rows = synth.frame_bgr_bottomup("photo", w, h, synth.SEED_BASE + 1)
    bmp = tmp_path / "img.bmp"
    bmp.write_bytes(oracle.bmp_wrap(rows, w, h))
    cli = os.path.join(host, "awi_cli")
    r = subprocess.run([cli, "convert", str(bmp)], capture_output=True, text=True)
    assert r.returncode == 0 and "convert completed!" in r.stdout, r.stderr
    awi = (tmp_path / "img.awi").read_bytes()
    want = port.encode_bgr(rows, w, h, synth.bmp_stride(w))
    assert awi == oracle.awi_header(w, h) + want.tobytes()
    os.remove(bmp)
    r = subprocess.run([cli, "revert", str(tmp_path / "img.awi")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    back = (tmp_path / "img.bmp").read_bytes()
    assert back[:54] == oracle.bmp_wrap(rows, w, h)[:54]
    assert back[54:] == port.decode(want, w, h)[0].tobytes()
    r = subprocess.run([cli, "show", str(tmp_path / "img.awi")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert (tmp_path / "img.rgb").read_bytes() == port.decode(want, w, h, rgb_topdown=True)[0].tobytes()
    # exit codes of main.c:72-86
    assert subprocess.run([cli, "convert", str(tmp_path / "img.awi")], capture_output=True).returncode == 2
    assert subprocess.run([cli, "revert", str(tmp_path / "nope.awi")], capture_output=True).returncode == 1


class Block16(C.Structure):
    _fields_ = [("b", (C.c_uint8 * 16) * 16)]


class RGBBlock16(C.Structure):
    _fields_ = [("r", Block16), ("g", Block16), ("b", Block16)]


class PixelInput(C.Structure):
    pass


class BitsOutput(C.Structure):
    pass


PixelInput._fields_ = [("context", C.c_void_p), ("width", C.c_int), ("height", C.c_int),
                       ("begin", C.CFUNCTYPE(C.c_bool, C.POINTER(PixelInput))),
                       ("read_block", C.CFUNCTYPE(C.c_bool, C.POINTER(PixelInput), C.c_int, C.c_int, C.POINTER(RGBBlock16))),
                       ("end", C.CFUNCTYPE(C.c_bool, C.POINTER(PixelInput)))]
BitsOutput._fields_ = [("context", C.c_void_p), ("buffer", C.c_uint8), ("bit_count", C.c_int),
                       ("write", C.CFUNCTYPE(C.c_bool, C.POINTER(BitsOutput), C.c_void_p, C.c_size_t))]


def test_compress_to_awi_through_plugin_vtables(host, port):
    """An arbitrary PixelInput/BitsOutput pair (here: Python callbacks) like any plug-in of pixelrw.h."""
    lib = C.CDLL(os.path.join(host, "libawi_gpu_host.so"))
    lib.compress_to_awi.restype = C.c_bool
    w, h = 70, 45
    rgb = synth.frame_rgb_topdown("noise", w, h, 99)
    order, sink = [], bytearray()

    def begin(self):
        self.contents.width, self.contents.height = w, h
        return True

    def read_block(self, bx, by, out):
        order.append((by, bx))
        blk = out.contents
        for name, ch in (("r", 0), ("g", 1), ("b", 2)):
            plane = np.zeros((16, 16), np.uint8)
            sub = rgb[16 * by:16 * by + 16, 16 * bx:16 * bx + 16, ch]
            plane[:sub.shape[0], :sub.shape[1]] = sub
            C.memmove(getattr(blk, name).b, plane.ctypes.data, 256)
        return True

    def write(self, data, size):
        sink.extend(C.string_at(data, size))
        return True

    pi = PixelInput(None, 0, 0, PixelInput._fields_[3][1](begin), PixelInput._fields_[4][1](read_block),
                    PixelInput._fields_[5][1](0))
    bo = BitsOutput(None, 0, 0, BitsOutput._fields_[3][1](write))
    assert lib.compress_to_awi(C.byref(pi), C.byref(bo))
    assert order == [(y, x) for y in range(3) for x in range(5)]            # strict raster order
    rows = synth.rows_from_rgb_topdown(rgb)
    want = port.encode_bgr(rows, w, h, synth.bmp_stride(w))
    assert bytes(sink) == oracle.awi_header(w, h) + want.tobytes()


def test_reference_bmp_code_linked_against_our_awi(port, tmp_path):
    """oracle/_ref/ref_gpu_cli = the reference's OWN bmp.c / bitrw.c / utils.c and headers linked with
    host/awi_gpu.c in place of the reference's awi.c (built where /root/reference exists; the binary
    travels).  The reference's BMP reader drives our compress_to_awi block by block, our decompress_awi
    drives the reference's BMP writer."""
    cli = os.path.join(ROOT, "oracle", "_ref", "ref_gpu_cli")
    if not os.path.exists(cli):
        pytest.skip("oracle/_ref/ref_gpu_cli not built on this machine")
    w, h = 332, 205                                  # 3w is a multiple of 4: the reference's row padding is indeterminate
    rows = synth.frame_bgr_bottomup("photo", w, h, synth.SEED_BASE + 2)
    bmp = tmp_path / "in.bmp"
    # pixel array at offset 138 like the reference's own sample (bfOffBits is honoured, bmp.c:35)
    body = oracle.bmp_wrap(rows, w, h)
    import struct
    bmp.write_bytes(body[:10] + struct.pack("<I", 138) + body[14:54] + b"\0" * 84 + body[54:])
    r = subprocess.run([cli, "convert", str(bmp), str(tmp_path / "out.awi")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    want = port.encode_bgr(rows, w, h, synth.bmp_stride(w))
    assert (tmp_path / "out.awi").read_bytes() == oracle.awi_header(w, h) + want.tobytes()
    r = subprocess.run([cli, "revert", str(tmp_path / "out.awi"), str(tmp_path / "back.bmp")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    back = (tmp_path / "back.bmp").read_bytes()
    assert back[54:] == port.decode(want, w, h)[0].tobytes()


def test_repeated_calls_are_deterministic(port):
    """Back-to-back calls reuse everything a context owns (staging cursor, chained offsets, page-locked slots,
    result blocks): the same frames must give the same bytes every time, through every host entry point,
    from page-locked and from ordinary memory, with other sizes in between."""
    import ctypes as C
    import torch
    from image_compression_b200 import codec
    codec.init([0])
    lib = codec.load()
    w, h = 1920, 1080
    stride = synth.bmp_stride(w)
    frames = [synth.frame_bgr_bottomup(k, w, h, 0xA11CE08, f).reshape(-1) for f, k in enumerate(("photo", "noise", "photo"))]
    want = [port.encode_bgr(f, w, h, stride).tobytes() for f in frames]
    small = synth.frame_bgr_bottomup("noise", 200, 120, 0xA11CE09).reshape(-1)
    want_small = port.encode_bgr(small, 200, 120, synth.bmp_stride(200)).tobytes()
    hp = lib.gpu_codec_host_alloc(frames[0].size)
    pinned = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_uint8)), shape=(frames[0].size,))
    try:
        for it in range(12):
            f = it % 3
            if it % 2:
                pinned[:] = frames[f]
                assert codec.encode_bgr(pinned, w, h, stride) == want[f], it
            else:
                assert codec.encode_bgr(frames[f], w, h, stride) == want[f], it
            if it % 4 == 1:
                assert codec.encode_bgr(small, 200, 120, synth.bmp_stride(200)) == want_small
            if it % 4 == 2:
                got = codec.encode_batch_bgr(frames, w, h, stride)
                assert [bytes(g) for g in got] == want
            if it % 4 == 3:
                px = codec.decode_to_bgr(np.frombuffer(want[f], np.uint8), w, h, stride)
                assert px.tobytes() == port.decode(np.frombuffer(want[f], np.uint8), w, h, stride)[0].tobytes()
    finally:
        lib.gpu_codec_host_free(C.c_void_p(hp))
