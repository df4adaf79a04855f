"""oracle — TEST INFRASTRUCTURE ONLY (never imported by image_compression_b200).

Two CPU checkers for the BMP->AWI hot path, both behind ctypes:

* ``oracle.port``  — ``liboracle.so``: the C restatement in ``awi_oracle.c``.
* ``oracle.ref``   — ``_ref/libawi_ref.so``: the UNMODIFIED reference sources compiled
  from ``/root/reference/src`` (``oracle/Makefile``) behind ``ref_shim.c``; ``None``
  when that build is absent.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this package.
"""
from __future__ import annotations

import ctypes as C
import os
import struct
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

_u8p = C.POINTER(C.c_uint8)
_i8p = C.POINTER(C.c_int8)
_i32p = C.POINTER(C.c_int)
_u32p = C.POINTER(C.c_uint32)


def _ptr(a: np.ndarray, typ):
    return a.ctypes.data_as(typ)


def build(force: bool = False) -> None:
    """Compile liboracle.so (and _ref when /root/reference exists). Building the
    checker is not using it."""
    so = os.path.join(_HERE, "liboracle.so")
    ref_so = os.path.join(_HERE, "_ref", "libawi_ref.so")
    need = force or not os.path.exists(so) or (
        os.path.getmtime(so) < os.path.getmtime(os.path.join(_HERE, "awi_oracle.c")))
    if os.path.isdir("/root/reference/src"):
        cli = os.path.join(_HERE, "_ref", "ref_gpu_cli")
        gpulib = os.path.join(os.path.dirname(_HERE), "image_compression_b200", "libgpu_codec.so")
        if not os.path.exists(ref_so) or (os.path.exists(gpulib) and not os.path.exists(cli)):
            need = True
    if need:
        subprocess.run(["make", "-C", _HERE], check=True, stdout=subprocess.DEVNULL)


def bmp_wrap(rows: np.ndarray, w: int, h: int) -> bytes:
    """54-byte BITMAPINFOHEADER file around bottom-up BGR rows (stride must be the
    BMP stride ((24w+31)/32)*4, bmp.c:36)."""
    stride = ((24 * w + 31) // 32) * 4
    rows = np.ascontiguousarray(rows, dtype=np.uint8).reshape(-1)
    assert rows.size == stride * h, (rows.size, stride, h)
    hdr = struct.pack("<2sIHHI", b"BM", 54 + rows.size, 0, 0, 54)
    hdr += struct.pack("<IiiHHIIiiII", 40, w, h, 1, 24, 0, rows.size, 2835, 2835, 0, 0)
    return hdr + rows.tobytes()


def awi_header(w: int, h: int) -> bytes:
    """awi.c:75-83 / awi.h:27-31."""
    return b"AWESOMEI" + struct.pack("<II", w, h)


class Port:
    """ctypes face of awi_oracle.c."""

    def __init__(self):
        build()
        self.lib = C.CDLL(os.path.join(_HERE, "liboracle.so"))
        L = self.lib
        L.awi_oracle_encode_rows.restype = C.c_size_t
        L.awi_oracle_encode_rows.argtypes = [_u8p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int,
                                             _u8p, C.c_size_t, _i8p, _u32p]
        L.awi_oracle_decode.restype = C.c_long
        L.awi_oracle_decode.argtypes = [_u8p, C.c_size_t, C.c_int, C.c_int, _u8p, C.c_size_t, C.c_int, _u32p]
        L.awi_oracle_decode_rows.restype = C.c_long
        L.awi_oracle_decode_rows.argtypes = [_u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t, _i32p,
                                             C.c_void_p, C.c_size_t, C.c_int, _u32p]
        L.awi_oracle_rle.restype = C.c_int
        L.awi_oracle_rle.argtypes = [_i32p, C.c_int, _u8p]
        L.awi_oracle_unrle.restype = C.c_int
        L.awi_oracle_unrle.argtypes = [_u8p, C.c_size_t, C.c_int, _i32p]
        for name in ("awi_oracle_transform",):
            getattr(L, name).argtypes = [_u8p, C.c_int, _i32p]
        L.awi_oracle_inverse_transform.argtypes = [_i32p, C.c_int, _u8p]
        L.awi_oracle_fdct.argtypes = [_u8p, _i32p]
        L.awi_oracle_idct.argtypes = [_i32p, _u8p]
        L.awi_oracle_quant.argtypes = [_i32p, C.c_int, _i32p]
        L.awi_oracle_rgb2ycc_n.argtypes = [_u8p, C.c_size_t, _u8p]
        L.awi_oracle_ycc2rgb_n.argtypes = [_u8p, C.c_size_t, _u8p]
        L.awi_oracle_cos_table.argtypes = [C.POINTER(C.c_float)]
        L.awi_oracle_tables.argtypes = [_i32p, _i32p, _i32p]

    # -- frame paths -------------------------------------------------------
    def encode_rows(self, rows: np.ndarray, w: int, h: int, stride: int, mb_row0: int, mb_row1: int,
                    want_coefs: bool = False):
        rows = np.ascontiguousarray(rows, dtype=np.uint8).reshape(-1)
        mbw = (w + 15) // 16
        nch = (mb_row1 - mb_row0) * mbw * 6
        cap = nch * 121 + 16
        out = np.empty(cap, np.uint8)
        q = np.empty((nch, 64), np.int8) if want_coefs else None
        ln = np.empty(nch, np.uint32) if want_coefs else None
        n = self.lib.awi_oracle_encode_rows(_ptr(rows, _u8p), w, h, stride, mb_row0, mb_row1,
                                            _ptr(out, _u8p), cap,
                                            _ptr(q, _i8p) if want_coefs else None,
                                            _ptr(ln, _u32p) if want_coefs else None)
        assert n != C.c_size_t(-1).value
        payload = out[:n].copy()
        return (payload, q, ln) if want_coefs else payload

    def encode_bgr(self, rows, w, h, stride, want_coefs=False):
        return self.encode_rows(rows, w, h, stride, 0, (h + 15) // 16, want_coefs)

    def decode(self, payload: np.ndarray, w: int, h: int, stride: int | None = None, rgb_topdown=False,
               want_offsets=False):
        payload = np.ascontiguousarray(payload, dtype=np.uint8).reshape(-1)
        if stride is None:
            stride = ((24 * w + 31) // 32) * 4
        dst = np.zeros(w * h * 3 if rgb_topdown else stride * h, np.uint8)
        nch = ((w + 15) // 16) * ((h + 15) // 16) * 6
        off = np.empty(nch, np.uint32) if want_offsets else None
        used = self.lib.awi_oracle_decode(_ptr(payload, _u8p), payload.size, w, h, _ptr(dst, _u8p), stride,
                                          1 if rgb_topdown else 0, _ptr(off, _u32p) if want_offsets else None)
        if used < 0:
            raise ValueError("malformed AWI payload")
        return (dst, used, off) if want_offsets else (dst, used)

    def decode_rows(self, payload: np.ndarray, w: int, h: int, mb_row0: int, mb_row1: int, start: int, pred,
                    dst: np.ndarray, stride: int, rgb_topdown=False) -> int:
        """Macroblock rows [mb_row0, mb_row1) from byte `start` with the six predictors `pred` as they stand
        there, into the full-frame buffer `dst` (only the band is written).  Returns the offset after the band."""
        pr = np.ascontiguousarray(pred, np.int32)
        end = self.lib.awi_oracle_decode_rows(_ptr(payload, _u8p), payload.size, w, h, mb_row0, mb_row1, start,
                                              _ptr(pr, _i32p), dst.ctypes.data, stride, 1 if rgb_topdown else 0, None)
        if end < 0:
            raise ValueError("malformed AWI payload")
        return int(end)

    # -- stages ------------------------------------------------------------
    def transform(self, px: np.ndarray, chroma: int) -> np.ndarray:
        px = np.ascontiguousarray(px, np.uint8).reshape(64)
        out = np.empty(64, np.int32)
        self.lib.awi_oracle_transform(_ptr(px, _u8p), chroma, _ptr(out, _i32p))
        return out

    def fdct(self, px):
        px = np.ascontiguousarray(px, np.uint8).reshape(64)
        out = np.empty(64, np.int32)
        self.lib.awi_oracle_fdct(_ptr(px, _u8p), _ptr(out, _i32p))
        return out

    def quant(self, coef, chroma):
        coef = np.ascontiguousarray(coef, np.int32).reshape(64)
        out = np.empty(64, np.int32)
        self.lib.awi_oracle_quant(_ptr(coef, _i32p), chroma, _ptr(out, _i32p))
        return out

    def idct(self, coef):
        coef = np.ascontiguousarray(coef, np.int32).reshape(64)
        out = np.empty(64, np.uint8)
        self.lib.awi_oracle_idct(_ptr(coef, _i32p), _ptr(out, _u8p))
        return out

    def inverse_transform(self, chunk, chroma):
        chunk = np.ascontiguousarray(chunk, np.int32).reshape(64)
        out = np.empty(64, np.uint8)
        self.lib.awi_oracle_inverse_transform(_ptr(chunk, _i32p), chroma, _ptr(out, _u8p))
        return out

    def rle(self, chunk, prev_dc) -> bytes:
        chunk = np.ascontiguousarray(chunk, np.int32).reshape(64)
        out = np.empty(160, np.uint8)
        n = self.lib.awi_oracle_rle(_ptr(chunk, _i32p), prev_dc, _ptr(out, _u8p))
        return out[:n].tobytes()

    def unrle(self, data: bytes, prev_dc):
        buf = np.frombuffer(data, np.uint8).copy()
        out = np.empty(64, np.int32)
        used = self.lib.awi_oracle_unrle(_ptr(buf, _u8p), buf.size, prev_dc, _ptr(out, _i32p))
        return out, used

    def rgb2ycc(self, rgb):
        rgb = np.ascontiguousarray(rgb, np.uint8).reshape(-1, 3)
        out = np.empty_like(rgb)
        self.lib.awi_oracle_rgb2ycc_n(_ptr(rgb, _u8p), rgb.shape[0], _ptr(out, _u8p))
        return out

    def ycc2rgb(self, ycc):
        ycc = np.ascontiguousarray(ycc, np.uint8).reshape(-1, 3)
        out = np.empty_like(ycc)
        self.lib.awi_oracle_ycc2rgb_n(_ptr(ycc, _u8p), ycc.shape[0], _ptr(out, _u8p))
        return out

    def cos_table(self):
        t = np.empty(64, np.float32)
        self.lib.awi_oracle_cos_table(_ptr(t, C.POINTER(C.c_float)))
        return t.reshape(8, 8)

    def tables(self):
        a, b, c = (np.empty(64, np.int32) for _ in range(3))
        self.lib.awi_oracle_tables(_ptr(a, _i32p), _ptr(b, _i32p), _ptr(c, _i32p))
        return a, b, c


class Ref:
    """ctypes face of the unmodified reference build (oracle/_ref)."""

    def __init__(self, variant: str = ""):
        path = os.path.join(_HERE, "_ref", f"libawi_ref{variant}.so")
        if not os.path.exists(path):
            build()
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = C.CDLL(path)
        L = self.lib
        L.ref_encode_bmp_mem.restype = C.c_size_t
        L.ref_encode_bmp_mem.argtypes = [_u8p, C.c_size_t, _u8p, C.c_size_t]
        L.ref_decode_awi_to_bmp_mem.restype = C.c_size_t
        L.ref_decode_awi_to_bmp_mem.argtypes = [_u8p, C.c_size_t, _u8p, C.c_size_t]
        L.ref_decode_awi_to_rgb_mem.restype = C.c_int
        L.ref_decode_awi_to_rgb_mem.argtypes = [_u8p, C.c_size_t, _u8p, C.c_size_t, _i32p, _i32p]
        L.ref_transform_block.argtypes = [_u8p, C.c_int, _i32p]
        L.ref_dct_block.argtypes = [_u8p, _i32p]
        L.ref_quantize_block.argtypes = [_i32p, C.c_int, _i32p]
        L.ref_inverse_transform_block.argtypes = [_i32p, C.c_int, _u8p]
        L.ref_rle_encode_chunk.restype = C.c_int
        L.ref_rle_encode_chunk.argtypes = [_i32p, C.c_int, _u8p, C.c_size_t]
        L.ref_rle_decode_chunk.restype = C.c_int
        L.ref_rle_decode_chunk.argtypes = [_u8p, C.c_size_t, C.c_int, _i32p]
        L.ref_rgb_to_ycbcr_px.argtypes = [_u8p, C.c_size_t, _u8p]
        L.ref_ycbcr_to_rgb_px.argtypes = [_u8p, C.c_size_t, _u8p]
        L.ref_warm()

    def encode_bmp(self, bmp: bytes | np.ndarray) -> bytes:
        buf = np.frombuffer(bmp, np.uint8) if isinstance(bmp, (bytes, bytearray)) else bmp
        buf = np.ascontiguousarray(buf)
        w, h = struct.unpack_from("<ii", buf[:26].tobytes(), 18)
        cap = 16 + ((w + 15) // 16) * ((h + 15) // 16) * 6 * 121
        out = np.empty(cap, np.uint8)
        n = self.lib.ref_encode_bmp_mem(_ptr(buf, _u8p), buf.size, _ptr(out, _u8p), cap)
        if n == 0:
            raise RuntimeError("reference compress_to_awi failed")
        return out[:n].tobytes()

    def encode_bgr(self, rows, w, h) -> np.ndarray:
        """payload only (header stripped)."""
        awi = self.encode_bmp(bmp_wrap(rows, w, h))
        return np.frombuffer(awi, np.uint8)[16:].copy()

    def decode_to_bmp(self, awi: bytes) -> bytes:
        buf = np.frombuffer(awi, np.uint8).copy()
        w, h = struct.unpack_from("<II", awi, 8)
        cap = 54 + ((24 * w + 31) // 32) * 4 * h
        out = np.empty(cap, np.uint8)
        n = self.lib.ref_decode_awi_to_bmp_mem(_ptr(buf, _u8p), buf.size, _ptr(out, _u8p), cap)
        if n == 0:
            raise RuntimeError("reference decompress_awi failed")
        return out[:n].tobytes()

    def decode_to_rgb(self, awi: bytes):
        buf = np.frombuffer(awi, np.uint8).copy()
        w, h = struct.unpack_from("<II", awi, 8)
        out = np.empty(w * h * 3, np.uint8)
        ww, hh = C.c_int(), C.c_int()
        ok = self.lib.ref_decode_awi_to_rgb_mem(_ptr(buf, _u8p), buf.size, _ptr(out, _u8p), out.size,
                                                C.byref(ww), C.byref(hh))
        if not ok:
            raise RuntimeError("reference decompress_awi failed")
        return out.reshape(h, w, 3)

    def transform(self, px, chroma):
        px = np.ascontiguousarray(px, np.uint8).reshape(64)
        out = np.empty(64, np.int32)
        self.lib.ref_transform_block(_ptr(px, _u8p), chroma, _ptr(out, _i32p))
        return out

    def fdct(self, px):
        px = np.ascontiguousarray(px, np.uint8).reshape(64)
        out = np.empty(64, np.int32)
        self.lib.ref_dct_block(_ptr(px, _u8p), _ptr(out, _i32p))
        return out

    def quant(self, coef, chroma):
        coef = np.ascontiguousarray(coef, np.int32).reshape(64)
        out = np.empty(64, np.int32)
        self.lib.ref_quantize_block(_ptr(coef, _i32p), chroma, _ptr(out, _i32p))
        return out

    def inverse_transform(self, chunk, chroma):
        chunk = np.ascontiguousarray(chunk, np.int32).reshape(64)
        out = np.empty(64, np.uint8)
        self.lib.ref_inverse_transform_block(_ptr(chunk, _i32p), chroma, _ptr(out, _u8p))
        return out

    def rle(self, chunk, prev_dc) -> bytes:
        chunk = np.ascontiguousarray(chunk, np.int32).reshape(64)
        out = np.empty(256, np.uint8)
        n = self.lib.ref_rle_encode_chunk(_ptr(chunk, _i32p), prev_dc, _ptr(out, _u8p), out.size)
        assert n >= 0
        return out[:n].tobytes()

    def unrle(self, data: bytes, prev_dc):
        buf = np.frombuffer(data, np.uint8).copy()
        out = np.empty(64, np.int32)
        used = self.lib.ref_rle_decode_chunk(_ptr(buf, _u8p), buf.size, prev_dc, _ptr(out, _i32p))
        return out, used

    def rgb2ycc(self, rgb):
        rgb = np.ascontiguousarray(rgb, np.uint8).reshape(-1, 3)
        out = np.empty_like(rgb)
        self.lib.ref_rgb_to_ycbcr_px(_ptr(rgb, _u8p), rgb.shape[0], _ptr(out, _u8p))
        return out

    def ycc2rgb(self, ycc):
        ycc = np.ascontiguousarray(ycc, np.uint8).reshape(-1, 3)
        out = np.empty_like(ycc)
        self.lib.ref_ycbcr_to_rgb_px(_ptr(ycc, _u8p), ycc.shape[0], _ptr(out, _u8p))
        return out


_port = None
_ref = {}


def port() -> Port:
    global _port
    if _port is None:
        _port = Port()
    return _port


def ref(variant: str = "") -> Ref | None:
    if variant not in _ref:
        try:
            _ref[variant] = Ref(variant)
        except (FileNotFoundError, OSError):
            _ref[variant] = None
    return _ref[variant]
