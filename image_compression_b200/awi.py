"""AWI container + codec drivers — mirror of awi.c / awi.h.

`compress_to_awi(pi, bo)` and `decompress_awi(bi, po)` keep the reference's signatures and
error behaviour (bool result, one stderr line) — awi.h:45-46.  Their macroblock loops
(awi.c:94-128, 184-217) are one call into the CUDA library each; header I/O stays here
(awi.c:75-83, 155-165)."""
from __future__ import annotations

import struct
import sys
from typing import BinaryIO

import numpy as np

from . import codec
from .bmp import bmp_row_size
from .pixelrw import PixelInput, PixelOutput, RGBBlock16

MAGIC = b"AWESOMEI"                     # awi.c:76
HEADER_BYTES = 16                        # awi.h:27-31


class BitsOutput:
    """bitrw.h:10-21: a byte sink taking arbitrary-size writes."""

    def write(self, data: bytes) -> bool:
        raise NotImplementedError


class BitsInput:
    """bitrw.h:32-43."""

    def read(self, size: int) -> bytes:
        raise NotImplementedError


class AWIFileOutput(BitsOutput):
    """init_awi_bits_output + awi_output_write (awi.c:18-33)."""

    def __init__(self, fp: BinaryIO):
        self.fp = fp

    def write(self, data: bytes) -> bool:
        self.fp.write(data)
        return True


class AWIFileInput(BitsInput):
    """init_awi_bits_input + awi_input_read (awi.c:38-53)."""

    def __init__(self, fp: BinaryIO):
        self.fp = fp

    def read(self, size: int) -> bytes:
        return self.fp.read(size) if size >= 0 else self.fp.read()


def pack_header(width: int, height: int) -> bytes:
    return MAGIC + struct.pack("<II", width, height)


def _gather_frame(pi: PixelInput):
    """Slow path for arbitrary plug-in sources: read_block in raster order (pixelrw.h) into
    bottom-up BGR rows."""
    w, h = pi.width, pi.height
    stride = bmp_row_size(w)
    img = np.zeros((h, stride), np.uint8)
    blk = RGBBlock16()
    for by in range((h + 15) // 16):
        for bx in range((w + 15) // 16):
            if not pi.read_block(bx, by, blk):
                return None, 0
            y0, x0 = 16 * by, 16 * bx
            hh, ww = min(16, h - y0), min(16, w - x0)
            rows = h - 1 - np.arange(y0, y0 + hh)
            img[rows, 3 * x0: 3 * (x0 + ww)] = np.stack(
                [blk.b[:hh, :ww], blk.g[:hh, :ww], blk.r[:hh, :ww]], -1).reshape(hh, 3 * ww)
    return img.reshape(-1), stride


def compress_to_awi(pi: PixelInput, bo: BitsOutput) -> bool:
    """awi.c:65-141."""
    try:
        if not pi.begin():
            print("Compressing: Preprocessing error!", file=sys.stderr)
            return False
        if not bo.write(pack_header(pi.width, pi.height)):
            print("Writing header error!", file=sys.stderr)
            return False
        bulk = pi.rows_bgr_bottomup()
        if bulk is None:
            rows, stride = _gather_frame(pi)
            if rows is None:
                print("Reading RGB blocks error!", file=sys.stderr)
                return False
        else:
            rows, stride = bulk
        payload = codec.encode_bgr(rows, pi.width, pi.height, stride)
        if not bo.write(payload):
            print("Encoding and writing error!", file=sys.stderr)
            return False
        if not pi.end():
            print("Compressing: Postprocessing error!", file=sys.stderr)
            return False
        return True
    except (IOError, codec.GpuCodecError) as e:
        print(f"Encoding and writing error! ({e})", file=sys.stderr)
        return False


def decompress_awi(bi: BitsInput, po: PixelOutput) -> bool:
    """awi.c:153-230."""
    try:
        head = bi.read(HEADER_BYTES)
        if len(head) != HEADER_BYTES:
            print("Reading header error!", file=sys.stderr)
            return False
        if head[:8] != MAGIC:
            print("Invalid file format!", file=sys.stderr)
            return False
        width, height = struct.unpack_from("<II", head, 8)
        if not po.begin(width, height):
            print("Decompressing: Preprocessing error!", file=sys.stderr)
            return False
        payload = bi.read(-1)
        if po.bulk_layout == "bgr_bottomup":
            stride = bmp_row_size(width)
            px = codec.decode_to_bgr(payload, width, height, stride)
            ok = po.write_frame(px, stride)
        elif po.bulk_layout == "rgb_topdown":
            ok = po.write_frame(codec.decode_to_rgb24(payload, width, height), 3 * width)
        else:   # arbitrary sink: replay write_block in raster order
            rgb = codec.decode_to_rgb24(payload, width, height)
            blk = RGBBlock16()
            ok = True
            for by in range((height + 15) // 16):
                for bx in range((width + 15) // 16):
                    y0, x0 = 16 * by, 16 * bx
                    hh, ww = min(16, height - y0), min(16, width - x0)
                    for i, plane in enumerate((blk.r, blk.g, blk.b)):
                        plane[:] = 0
                        plane[:hh, :ww] = rgb[y0:y0 + hh, x0:x0 + ww, i]
                    ok = ok and po.write_block(bx, by, blk)
        if not ok:
            print("Writing rgb blocks error!", file=sys.stderr)
            return False
        if not po.end():
            print("Decompressing: Postprocessing error!", file=sys.stderr)
            return False
        return True
    except (IOError, codec.GpuCodecError) as e:
        print(f"Reading and decoding error! ({e})", file=sys.stderr)
        return False
