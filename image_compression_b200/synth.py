"""Integer-only synthetic frames — identical in numpy and in the CUDA
generator (`gpu_codec_synth_bgr_dev`), so every rank / the CPU baseline / the tests can
rebuild the same bytes from (kind, seed, frame index) without shipping pixels around.

Layout is what the reference's BMP reader consumes (bmp.c:35-36,80-81,111-113):
bottom-up rows of B,G,R bytes, `stride = ((24w+31)/32)*4` bytes apart, padding bytes 0.
"""
from __future__ import annotations

import numpy as np

KINDS = {"noise": 0, "photo": 1, "flat": 2}
SEED_BASE = 0x0A11CE00


def bmp_stride(w: int) -> int:
    return ((24 * w + 31) // 32) * 4


def _fmix32(h: np.ndarray) -> np.ndarray:
    h = h.astype(np.uint32, copy=True)
    h ^= h >> np.uint32(16)
    h *= np.uint32(0x85EBCA6B)
    h ^= h >> np.uint32(13)
    h *= np.uint32(0xC2B2AE35)
    h ^= h >> np.uint32(16)
    return h


def _tri(t: np.ndarray) -> np.ndarray:
    return np.abs((t & 511) - 256) - 128


def frame_rgb_topdown(kind: str, w: int, h: int, seed: int, frame: int = 0) -> np.ndarray:
    """(h, w, 3) uint8, top-down, channel order R,G,B."""
    if kind == "flat":
        return np.full((h, w, 3), 144, np.uint8)
    s = np.uint32((seed ^ ((frame * 0x9E3779B9) & 0xFFFFFFFF)) & 0xFFFFFFFF)
    with np.errstate(over="ignore"):
        y = np.arange(h, dtype=np.uint32)[:, None, None]
        x = np.arange(w, dtype=np.uint32)[None, :, None]
        c = np.arange(3, dtype=np.uint32)[None, None, :]
        i = (y * np.uint32(w) + x) * np.uint32(3) + c
        m = _fmix32(i * np.uint32(0x9E3779B1) + s)
    if kind == "noise":
        return (m >> np.uint32(24)).astype(np.uint8)
    if kind == "photo":
        xi = x.astype(np.int32)
        yi = y.astype(np.int32)
        ci = c.astype(np.int32)
        grad = (_tri(3 * xi + 37 * ci + 5 * frame) + _tri(2 * yi + 91 * ci)) >> 1
        grain = (m >> np.uint32(27)).astype(np.int32) - 16
        return np.clip(128 + grad + grain, 0, 255).astype(np.uint8)
    raise ValueError(kind)


def frame_bgr_bottomup(kind: str, w: int, h: int, seed: int, frame: int = 0) -> np.ndarray:
    """(h, stride) uint8: row 0 is the BOTTOM image row, bytes B,G,R per pixel."""
    rgb = frame_rgb_topdown(kind, w, h, seed, frame)
    stride = bmp_stride(w)
    rows = np.zeros((h, stride), np.uint8)
    rows[:, : 3 * w] = rgb[::-1, :, ::-1].reshape(h, 3 * w)
    return rows


def rows_from_rgb_topdown(rgb: np.ndarray) -> np.ndarray:
    h, w, _ = rgb.shape
    rows = np.zeros((h, bmp_stride(w)), np.uint8)
    rows[:, : 3 * w] = rgb[::-1, :, ::-1].reshape(h, 3 * w)
    return rows
