"""Macroblock-row stripe / whole-frame batch partitioning and host-side stitching
(SURVEY.md §8e).  Pure host logic: no collective is needed because chunks are byte
aligned (rle.c:107) and the only cross-stripe state, six DC predictors (awi.c:91-92), is
recomputed by the kernel from one macroblock row of halo."""
from __future__ import annotations

from typing import List, Sequence, Tuple


def stripe_rows(mb_rows: int, n: int) -> List[Tuple[int, int]]:
    """Balanced split of mb_rows macroblock rows into n contiguous stripes [r0, r1)."""
    return [(mb_rows * g // n, mb_rows * (g + 1) // n) for g in range(n)]


def stripe_window(height: int, r0: int, r1: int) -> Tuple[int, int]:
    """Top-down pixel rows [y0, y1) a stripe needs: its own plus 16 rows of halo above."""
    return (max(0, 16 * r0 - 16), min(height, 16 * r1))


def window_byte_range(height: int, stride: int, y0: int, y1: int) -> Tuple[int, int]:
    """Byte range of rows y0..y1 inside a bottom-up pixel array (bmp.c:80-81)."""
    return ((height - y1) * stride, (height - y0) * stride)


def batch_share(n_frames: int, rank: int, world: int) -> range:
    """Frames a rank owns in an image-parallel batch (whole frames, round robin)."""
    return range(rank, n_frames, world)


def stitch(segments: Sequence[bytes]) -> bytes:
    """Concatenate per-stripe payload segments in stripe order: an exclusive prefix sum of
    their byte lengths gives each segment's offset — byte seams, no bit shifting."""
    return b"".join(segments)


def segment_offsets(lengths: Sequence[int]) -> List[int]:
    off, out = 0, []
    for n in lengths:
        out.append(off)
        off += n
    return out
