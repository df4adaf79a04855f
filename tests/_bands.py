"""Whole-frame oracle runs for the large configs, spread over host threads (ctypes releases the GIL).

The reference is single-threaded, and so are both checkers; a frame is cut into bands of macroblock
rows, which is legitimate because a band's bytes depend on what lies above it only through the six DC
predictors (awi.c:91-92), i.e. through the quantised DCs of the one macroblock in front of the band:

* port (oracle/awi_oracle.c): `awi_oracle_encode_rows` seeds the predictors from that macroblock itself;
  `awi_oracle_decode_rows` takes them as an argument (here: running sums of the 12-bit DC fields of the
  ORACLE's own payload, rle.c:118-128).
* ref (the unmodified reference build, where oracle/_ref travelled): each band is wrapped as a BMP of its own
  together with the macroblock row above it and encoded by the reference's compress_to_awi; the bytes of
  that extra row are dropped (they differ from the frame's only in the first macroblock's six DC fields,
  which are 12 bits wide whatever their value, so their length is the frame's).
"""
from __future__ import annotations

import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np


def _threads() -> int:
    return max(1, min(64, os.cpu_count() or 1))


def bands(mbh: int, rows_per_band: int = 16):
    return [(r, min(mbh, r + rows_per_band)) for r in range(0, mbh, rows_per_band)]


def port_encode_frame(port, rows: np.ndarray, w: int, h: int, stride: int, rows_per_band: int = 16):
    """(payload bytes, chunk byte lengths) of the whole frame by the C restatement."""
    mbh = (h + 15) // 16
    rows = np.ascontiguousarray(rows, np.uint8).reshape(-1)

    def one(b):
        return port.encode_rows(rows, w, h, stride, b[0], b[1], want_coefs=True)

    with ThreadPoolExecutor(_threads()) as ex:
        parts = list(ex.map(one, bands(mbh, rows_per_band)))
    payload = np.concatenate([p[0] for p in parts])
    lens = np.concatenate([p[2] for p in parts])
    return payload, lens


def ref_check_frame(ref, oracle_mod, rows: np.ndarray, w: int, h: int, stride: int, payload: np.ndarray, offs: np.ndarray,
                    rows_per_band: int = 16) -> None:
    """Every byte of `payload` (chunk offsets `offs`, one per chunk) against the unmodified reference."""
    assert stride == ((24 * w + 31) // 32) * 4
    mbw, mbh = (w + 15) // 16, (h + 15) // 16
    rows2 = np.ascontiguousarray(rows, np.uint8).reshape(h, stride)
    ends = np.concatenate([offs.astype(np.int64), [payload.size]])

    def one(b):
        r0, r1 = b
        ra = max(0, r0 - 1)                                   # one macroblock row of run-in for the predictors
        y0, y1 = 16 * ra, min(h, 16 * r1)
        # bottom-up storage: image rows y0..y1-1 are array rows h-y1 .. h-y0-1
        sub = rows2[h - y1: h - y0]
        hh = y1 - y0
        if y1 == h and h % 16 and ra > 0:
            pass                                              # a ragged last band keeps its ragged height: same zero padding
        got = ref.encode_bgr(sub, w, hh)
        skip = int(ends[r0 * mbw * 6] - ends[ra * mbw * 6])
        want = payload[ends[r0 * mbw * 6]: ends[r1 * mbw * 6]]
        return bool(got[skip:].tobytes() == want.tobytes())

    with ThreadPoolExecutor(_threads()) as ex:
        ok = list(ex.map(one, bands(mbh, rows_per_band)))
    bad = [b for b, o in zip(bands(mbh, rows_per_band), ok) if not o]
    assert not bad, f"bands differing from the reference: {bad[:8]}"


def dc_predictors(payload: np.ndarray, offs: np.ndarray) -> np.ndarray:
    """Absolute quantised DC of every chunk = running sum per stream (chunk index mod 6) of the 12-bit
    differences at the head of each chunk (rle.c:118-128).  Shape (n_mb, 6)."""
    o = offs.astype(np.int64)
    raw = (payload[o].astype(np.int32) << 4) | (payload[o + 1].astype(np.int32) >> 4)
    diff = np.where(raw > 2047, raw - 4095, raw)
    return np.cumsum(diff.reshape(-1, 6), axis=0, dtype=np.int64).astype(np.int32)


def port_decode_frame(port, payload: np.ndarray, lens: np.ndarray, w: int, h: int, stride: int, rgb_topdown: bool = False,
                      rows_per_band: int = 16) -> np.ndarray:
    """Pixels of the whole frame by the C restatement; `payload`/`lens` as port_encode_frame returned them."""
    mbw, mbh = (w + 15) // 16, (h + 15) // 16
    payload = np.ascontiguousarray(payload, np.uint8)
    offs = np.concatenate([[0], np.cumsum(lens.astype(np.int64))])
    dc = dc_predictors(payload, offs[:-1])
    dst = np.zeros(w * h * 3 if rgb_topdown else stride * h, np.uint8)

    def one(b):
        r0, r1 = b
        pred = dc[r0 * mbw - 1] if r0 > 0 else np.zeros(6, np.int32)
        end = port.decode_rows(payload, w, h, r0, r1, int(offs[r0 * mbw * 6]), pred, dst, stride, rgb_topdown)
        return end == int(offs[r1 * mbw * 6])

    with ThreadPoolExecutor(_threads()) as ex:
        ok = list(ex.map(one, bands(mbh, rows_per_band)))
    assert all(ok)
    return dst
