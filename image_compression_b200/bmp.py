"""BMP pixel source / sink and the packed-RGB sink — mirror of bmp.c / bmp.h.

Host-side I/O only; what matters to the hot path is the layout contract the reference's
reader defines (bmp.c:25-119): 14+40 byte headers, `bfOffBits` honoured, rows bottom-up,
B,G,R order, stride ((24w+31)/32)*4, pixels beyond width/height read as (0,0,0)."""
from __future__ import annotations

import struct
from typing import BinaryIO, Optional

import numpy as np

from .pixelrw import PixelInput, PixelOutput, RGBBlock16


def bmp_row_size(width: int) -> int:
    return ((24 * width + 31) // 32) * 4           # bmp.c:36


def read_bmp_header(fp: BinaryIO):
    """bmp.c:327-343: reads BMPHeader + BMPInfoHeader (bmp.h:10-31), checks only 'BM'."""
    head = fp.read(14)
    info = fp.read(40)
    if len(head) != 14 or len(info) != 40:
        raise IOError("Reading bmp header error!")
    bf_type, _size, _r1, _r2, off_bits = struct.unpack("<HIHHI", head)
    if bf_type != 0x4D42:
        raise IOError("Invalid BMP format!")
    _bi_size, width, height = struct.unpack_from("<Iii", info, 0)
    return off_bits, width, height


class BMPPixelInput(PixelInput):
    """init_bmp_pixel_input (bmp.c:18-23) + begin/read_block/end (bmp.c:25-128)."""

    def __init__(self, fp: BinaryIO):
        self.fp = fp
        self._rows: Optional[np.ndarray] = None
        self.row_size = 0

    def begin(self) -> bool:
        self.pixel_data_offset, self.width, self.height = read_bmp_header(self.fp)
        self.row_size = bmp_row_size(self.width)
        return True

    def rows_bgr_bottomup(self):
        if self._rows is None:
            self.fp.seek(self.pixel_data_offset)
            need = self.row_size * self.height
            data = self.fp.read(need)
            if len(data) != need:
                raise IOError("Reading RGB blocks error!")
            self._rows = np.frombuffer(data, np.uint8)
        return self._rows, self.row_size

    def read_block(self, bx: int, by: int, out: RGBBlock16) -> bool:
        rows, stride = self.rows_bgr_bottomup()
        img = rows.reshape(self.height, stride)
        for plane in (out.r, out.g, out.b):
            plane[:] = 0
        y0, x0 = 16 * by, 16 * bx
        hh = max(0, min(16, self.height - y0))
        ww = max(0, min(16, self.width - x0))
        if hh and ww:
            # top-down row y is file row height-1-y (bmp.c:80-81)
            sub = img[self.height - 1 - np.arange(y0, y0 + hh)][:, 3 * x0: 3 * (x0 + ww)].reshape(hh, ww, 3)
            out.b[:hh, :ww], out.g[:hh, :ww], out.r[:hh, :ww] = sub[..., 0], sub[..., 1], sub[..., 2]
        return True


class BMPPixelOutput(PixelOutput):
    """init_bmp_pixel_output (bmp.c:138-143) + begin/write_block/end (bmp.c:145-267)."""

    bulk_layout = "bgr_bottomup"

    def __init__(self, fp: BinaryIO):
        self.fp = fp
        self._buf: Optional[np.ndarray] = None

    def begin(self, width: int, height: int) -> bool:
        self.width, self.height = width, height
        self.row_size = bmp_row_size(width)
        img_size = self.row_size * height
        self.fp.write(struct.pack("<HIHHI", 0x4D42, 14 + 40 + img_size, 0, 0, 54))           # bmp.c:161-167
        self.fp.write(struct.pack("<IiiHHIIiiII", 40, width, height, 1, 24, 0, img_size, 2835, 2835, 0, 0))
        return True

    def write_frame(self, pixels: np.ndarray, row_stride: int) -> bool:
        assert row_stride == self.row_size
        self.fp.write(pixels.tobytes())
        return True

    def write_block(self, bx: int, by: int, block: RGBBlock16) -> bool:
        if self._buf is None:
            self._buf = np.zeros((self.height, self.row_size), np.uint8)
        y0, x0 = 16 * by, 16 * bx
        hh = max(0, min(16, self.height - y0))
        ww = max(0, min(16, self.width - x0))
        if hh and ww:
            rows = self.height - 1 - np.arange(y0, y0 + hh)
            view = np.stack([block.b[:hh, :ww], block.g[:hh, :ww], block.r[:hh, :ww]], -1).reshape(hh, 3 * ww)
            self._buf[rows, 3 * x0: 3 * (x0 + ww)] = view
        return True

    def end(self) -> bool:
        if self._buf is not None:
            self.fp.write(self._buf.tobytes())
            self._buf = None
        return True


class RGBPixelOutput(PixelOutput):
    """init_rgb_pixel_output (bmp.c:273-320): packed top-down RGB24, the `show` sink
    (main.c:162-181).  The window itself (SDL) is not part of this library."""

    bulk_layout = "rgb_topdown"

    def __init__(self):
        self.pixels: Optional[np.ndarray] = None

    def begin(self, width: int, height: int) -> bool:
        self.width, self.height = width, height
        self.pixels = np.zeros((height, width, 3), np.uint8)          # bmp.c:287
        return True

    def write_frame(self, pixels: np.ndarray, row_stride: int) -> bool:
        self.pixels = pixels.reshape(self.height, self.width, 3)
        return True

    def write_block(self, bx: int, by: int, block: RGBBlock16) -> bool:
        y0, x0 = 16 * by, 16 * bx
        hh = max(0, min(16, self.height - y0))
        ww = max(0, min(16, self.width - x0))
        if hh and ww:
            self.pixels[y0:y0 + hh, x0:x0 + ww, 0] = block.r[:hh, :ww]
            self.pixels[y0:y0 + hh, x0:x0 + ww, 1] = block.g[:hh, :ww]
            self.pixels[y0:y0 + hh, x0:x0 + ww, 2] = block.b[:hh, :ww]
        return True
