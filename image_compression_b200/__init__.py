"""image_compression_b200 — B200 (sm_100a) implementation of the BMP->AWI encode hot path of
du207/image_compression and its inverse, behind the reference's own interfaces.

  csrc/            hand-written CUDA kernels + the C ABI (include/gpu_codec.h) -> libgpu_codec.so
  codec.py         ctypes binding of that C ABI (no CPU fallback)
  awi.py           compress_to_awi / decompress_awi + AWI container  (awi.c)
  bmp.py           BMP pixel source/sink and the RGB24 sink           (bmp.c, pixelrw.h)
  stripes.py       macroblock-row stripe / batch partitioning and host-side stitching
  synth.py         integer-only synthetic frames for tests and benchmarks
"""
from . import synth  # noqa: F401

__all__ = ["awi", "bmp", "codec", "stripes", "synth", "build"]
