"""Pixel plug-in API — mirror of pixelrw.h:9-44.  A source/sink exchanges 16x16 macroblocks
(`RGBBlock16`, block.h:15-23: three planar uint8[16][16]) in strict raster order; `begin`
and `end` are optional.  Sources/sinks that can hand over the whole frame at once expose
the bulk methods, which compress_to_awi / decompress_awi prefer (one GPU call per frame)."""
from __future__ import annotations

import numpy as np


class RGBBlock16:
    """block.h:15-23."""

    __slots__ = ("r", "g", "b")

    def __init__(self):
        self.r = np.zeros((16, 16), np.uint8)
        self.g = np.zeros((16, 16), np.uint8)
        self.b = np.zeros((16, 16), np.uint8)


class PixelInput:
    """pixelrw.h:9-25.  `width`/`height` are filled by begin()."""

    width = 0
    height = 0

    def begin(self) -> bool:                      # optional in the reference (NULL-checked, awi.c:66)
        return True

    def read_block(self, bx: int, by: int, out: RGBBlock16) -> bool:
        raise NotImplementedError

    def end(self) -> bool:
        return True

    # bulk fast path: bottom-up BGR rows + stride, or None when the source cannot provide it
    def rows_bgr_bottomup(self):
        return None


class PixelOutput:
    """pixelrw.h:28-44."""

    width = 0
    height = 0

    def begin(self, width: int, height: int) -> bool:
        self.width, self.height = width, height
        return True

    def write_block(self, bx: int, by: int, block: RGBBlock16) -> bool:
        raise NotImplementedError

    def end(self) -> bool:
        return True

    # bulk fast path: which layout the sink wants ("bgr_bottomup" / "rgb_topdown" / None)
    bulk_layout = None

    def write_frame(self, pixels: np.ndarray, row_stride: int) -> bool:
        raise NotImplementedError
