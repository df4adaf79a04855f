"""ctypes binding of libgpu_codec.so (include/gpu_codec.h) — the only way Python reaches
the kernels.  There is no CPU fallback: a missing library or GPU raises."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# GPU_CODEC_LIB: experiments only (tools/build_variant.py builds alternative kernels next to the product library)
LIB_PATH = os.environ.get("GPU_CODEC_LIB") or os.path.join(_HERE, "libgpu_codec.so")

_u8p = C.POINTER(C.c_uint8)

# every symbol include/gpu_codec.h declares: (name, restype, argtypes)
SYMBOLS = [
    ("gpu_codec_device_count", C.c_int, []),
    ("gpu_codec_init", C.c_bool, [C.c_int]),
    ("gpu_codec_init_devices", C.c_bool, [C.POINTER(C.c_int), C.c_int]),
    ("gpu_codec_shutdown", None, []),
    ("gpu_codec_last_error", C.c_char_p, []),
    ("gpu_codec_max_payload_bytes", C.c_size_t, [C.c_int, C.c_int]),
    ("gpu_codec_host_alloc", C.c_void_p, [C.c_size_t]),
    ("gpu_codec_host_free", None, [C.c_void_p]),
    ("gpu_codec_free", None, [C.c_void_p]),
    ("gpu_codec_encode_bgr", C.c_bool, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p),
                                        C.POINTER(C.c_size_t)]),
    ("gpu_codec_encode_batch_bgr", C.c_bool, [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_int,
                                              C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    ("gpu_codec_decode_to_bgr", C.c_bool, [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_int]),
    ("gpu_codec_decode_to_rgb24", C.c_bool, [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p]),
    ("gpu_codec_encode_bgr_dev", C.c_bool, [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_longlong, C.c_int, C.c_int,
                                            C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p]),
    ("gpu_codec_encode_batch_bgr_dev", C.c_bool, [C.c_int, C.c_void_p, C.c_int, C.c_longlong, C.c_int, C.c_int, C.c_longlong,
                                                  C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]),
    ("gpu_codec_encode_bgr_dev_noalloc", C.c_bool, [C.c_void_p, C.c_int, C.c_int, C.c_longlong, C.c_void_p, C.c_size_t,
                                                    C.c_void_p, C.c_void_p, C.c_void_p]),
    ("gpu_codec_decode_dev", C.c_bool, [C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int,
                                        C.c_void_p, C.c_longlong, C.c_int, C.c_void_p, C.c_void_p]),
    ("gpu_codec_sync_check", C.c_bool, [C.c_int, C.c_void_p]),
    ("gpu_codec_synth_bgr_dev", C.c_bool, [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_longlong, C.c_int, C.c_uint32,
                                           C.c_int, C.c_void_p]),
    ("gpu_codec_selftest_tile_buffer", None, [C.c_int]),
    ("gpu_codec_selftest_colour_dev", C.c_bool, [C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p]),
]

_lib = None


class GpuCodecError(RuntimeError):
    pass


def load() -> C.CDLL:
    """dlopen the in-tree library and type every entry point.  Raises when it has not been
    built (python -m image_compression_b200.build) — never falls back to anything else."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GpuCodecError(f"{LIB_PATH} is missing: build it with `python -m image_compression_b200.build` "
                                "(there is no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, res, args in SYMBOLS:
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def _check(ok: bool, what: str) -> None:
    if not ok:
        raise GpuCodecError(f"{what}: {load().gpu_codec_last_error().decode()}")


_initialised: Optional[tuple] = None


def init(devices: Optional[Sequence[int]] = None) -> None:
    """gpu_codec_init_devices; idempotent for the same device list."""
    global _initialised
    lib = load()
    if devices is None:
        devices = [0]
    devices = tuple(int(d) for d in devices)
    if _initialised == devices:
        return
    if _initialised is not None:
        lib.gpu_codec_shutdown()
        _initialised = None
    arr = (C.c_int * len(devices))(*devices)
    _check(lib.gpu_codec_init_devices(arr, len(devices)), "gpu_codec_init_devices")
    _initialised = devices


def shutdown() -> None:
    global _initialised
    if _lib is not None:
        _lib.gpu_codec_shutdown()
    _initialised = None


def n_contexts() -> int:
    return len(_initialised) if _initialised else 0


# ------------------------------------------------------------------ host-buffer calls

def _np_u8(a) -> np.ndarray:
    if isinstance(a, (bytes, bytearray, memoryview)):
        return np.frombuffer(a, np.uint8)
    return np.ascontiguousarray(a, dtype=np.uint8)


def encode_bgr(rows, width: int, height: int, row_stride: int) -> bytes:
    """Payload (no header) of a frame given as bottom-up BGR rows in host memory."""
    lib = load()
    rows = _np_u8(rows).reshape(-1)
    if rows.size < row_stride * height:
        raise ValueError("rows buffer smaller than row_stride*height")
    out = C.c_void_p()
    n = C.c_size_t()
    _check(lib.gpu_codec_encode_bgr(rows.ctypes.data, width, height, row_stride, C.byref(out), C.byref(n)),
           "gpu_codec_encode_bgr")
    try:
        return C.string_at(out.value, n.value)
    finally:
        lib.gpu_codec_free(out)


def encode_bgr_ptr(host_ptr: int, width: int, height: int, row_stride: int):
    """Same, from a raw host address (e.g. pinned memory); returns (malloc'd address, length);
    release with free_payload()."""
    lib = load()
    out = C.c_void_p()
    n = C.c_size_t()
    _check(lib.gpu_codec_encode_bgr(host_ptr, width, height, row_stride, C.byref(out), C.byref(n)),
           "gpu_codec_encode_bgr")
    return out.value, n.value


def free_payload(addr: int) -> None:
    load().gpu_codec_free(C.c_void_p(addr))


def encode_batch_bgr(frames: Sequence[np.ndarray], width: int, height: int, row_stride: int) -> list:
    lib = load()
    frames = [_np_u8(f).reshape(-1) for f in frames]
    n = len(frames)
    ptrs = (C.c_void_p * n)(*[f.ctypes.data for f in frames])
    outs = (C.c_void_p * n)()
    lens = (C.c_size_t * n)()
    _check(lib.gpu_codec_encode_batch_bgr(ptrs, n, width, height, row_stride, outs, lens), "gpu_codec_encode_batch_bgr")
    res = []
    for i in range(n):
        res.append(C.string_at(outs[i], lens[i]))
        lib.gpu_codec_free(C.c_void_p(outs[i]))
    return res


def decode_to_bgr(payload, width: int, height: int, row_stride: Optional[int] = None, out: Optional[np.ndarray] = None):
    lib = load()
    payload = _np_u8(payload).reshape(-1)
    if row_stride is None:
        row_stride = ((24 * width + 31) // 32) * 4
    if out is None:
        out = np.zeros(row_stride * height, np.uint8)
    _check(lib.gpu_codec_decode_to_bgr(payload.ctypes.data, payload.size, width, height, out.ctypes.data, row_stride),
           "gpu_codec_decode_to_bgr")
    return out


def decode_to_rgb24(payload, width: int, height: int) -> np.ndarray:
    lib = load()
    payload = _np_u8(payload).reshape(-1)
    out = np.zeros((height, width, 3), np.uint8)
    _check(lib.gpu_codec_decode_to_rgb24(payload.ctypes.data, payload.size, width, height, out.ctypes.data),
           "gpu_codec_decode_to_rgb24")
    return out


# ------------------------------------------------------------------ device-resident calls (torch tensors)

def _stream_ptr(stream) -> Optional[int]:
    if stream is None:
        import torch
        return torch.cuda.current_stream().cuda_stream
    return int(stream)


def encode_dev(ctx: int, rows, width: int, height: int, row_stride: int, payload, payload_len,
               window=None, mb_rows=None, chunk_off=None, qcoef=None, stream=None) -> None:
    """Asynchronous encode of device-resident rows (torch uint8 tensors). `payload_len` is a
    1-element int64/uint64 tensor."""
    lib = load()
    mbh = (height + 15) // 16
    wy0, wy1 = window if window is not None else (0, height)
    r0, r1 = mb_rows if mb_rows is not None else (0, mbh)
    _check(lib.gpu_codec_encode_bgr_dev(ctx, rows.data_ptr(), width, height, row_stride, wy0, wy1, r0, r1,
                                        payload.data_ptr(), payload.numel(), payload_len.data_ptr(),
                                        chunk_off.data_ptr() if chunk_off is not None else None,
                                        qcoef.data_ptr() if qcoef is not None else None, _stream_ptr(stream)),
           "gpu_codec_encode_bgr_dev")


def encode_batch_dev(ctx: int, frames, n_frames: int, frame_stride: int, width: int, height: int, row_stride: int,
                     payload, frame_off, stream=None) -> None:
    """gpu_codec_encode_batch_bgr_dev: n_frames frames in one launch; frame_off is an int64[n_frames+1] tensor."""
    _check(load().gpu_codec_encode_batch_bgr_dev(ctx, frames.data_ptr(), n_frames, frame_stride, width, height, row_stride,
                                                 payload.data_ptr(), payload.numel(), frame_off.data_ptr(),
                                                 _stream_ptr(stream)), "gpu_codec_encode_batch_bgr_dev")


def encode_dev_noalloc(rows, width: int, height: int, row_stride: int, payload, payload_len, status, stream=None) -> None:
    """gpu_codec_encode_bgr_dev_noalloc: whole frame on the current device, scratch inside the library
    (no allocation; at most 2,097,152 macroblocks). `payload_len` int64[1], `status` int32[1] tensors."""
    _check(load().gpu_codec_encode_bgr_dev_noalloc(rows.data_ptr(), width, height, row_stride, payload.data_ptr(),
                                                   payload.numel(), payload_len.data_ptr(), status.data_ptr(),
                                                   _stream_ptr(stream)), "gpu_codec_encode_bgr_dev_noalloc")


def decode_dev(ctx: int, payload, length: int, width: int, height: int, dst, row_stride: int, rgb_topdown: bool = False,
               mb_rows=None, chunk_off=None, dst_offset_bytes: int = 0, stream=None) -> None:
    lib = load()
    mbh = (height + 15) // 16
    r0, r1 = mb_rows if mb_rows is not None else (0, mbh)
    _check(lib.gpu_codec_decode_dev(ctx, payload.data_ptr(), length, width, height, r0, r1,
                                    dst.data_ptr() + dst_offset_bytes, row_stride, 1 if rgb_topdown else 0,
                                    chunk_off.data_ptr() if chunk_off is not None else None, _stream_ptr(stream)),
           "gpu_codec_decode_dev")


def sync_check(ctx: int = 0, stream=None) -> None:
    _check(load().gpu_codec_sync_check(ctx, _stream_ptr(stream)), "gpu_codec_sync_check")


def synth_dev(ctx: int, rows, width: int, height: int, row_stride: int, kind: int, seed: int, frame: int = 0,
              stream=None) -> None:
    _check(load().gpu_codec_synth_bgr_dev(ctx, rows.data_ptr(), width, height, row_stride, kind, seed & 0xFFFFFFFF, frame,
                                          _stream_ptr(stream)), "gpu_codec_synth_bgr_dev")


def selftest_colour_dev(ctx: int, src, inverse: bool, dst, stream=None) -> None:
    _check(load().gpu_codec_selftest_colour_dev(ctx, src.data_ptr(), src.numel() // 3, 1 if inverse else 0,
                                                dst.data_ptr(), _stream_ptr(stream)), "gpu_codec_selftest_colour_dev")


def max_payload_bytes(width: int, height: int) -> int:
    return load().gpu_codec_max_payload_bytes(width, height)
