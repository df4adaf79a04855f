"""CPU: the C-ABI library builds, loads, and exports every symbol include/gpu_codec.h declares;
without a GPU every entry point fails loudly instead of falling back."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from image_compression_b200 import build, codec
    build.build()
    return codec.load()


def test_exports_every_declared_symbol(lib):
    from image_compression_b200 import codec
    hdr = open(os.path.join(ROOT, "include", "gpu_codec.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(gpu_codec_\w+)\s*\(", hdr))
    assert len(declared) >= 18
    bound = {name for name, _, _ in codec.SYMBOLS}
    assert declared == bound, declared ^ bound
    for name in declared:
        assert hasattr(lib, name), name


def test_fails_loudly_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from image_compression_b200 import codec
    assert lib.gpu_codec_device_count() == 0
    with pytest.raises(codec.GpuCodecError, match="no CPU path"):
        codec.init([0])
    out, n = C.c_void_p(), C.c_size_t()
    buf = (C.c_uint8 * 48)()
    assert lib.gpu_codec_encode_bgr(buf, 4, 4, 12, C.byref(out), C.byref(n)) is False
    assert b"not initialised" in lib.gpu_codec_last_error()
    assert lib.gpu_codec_decode_to_rgb24(buf, 18, 4, 4, buf) is False


def test_max_payload_bound(lib):
    assert lib.gpu_codec_max_payload_bytes(640, 427) == 40 * 27 * 6 * 121
    assert lib.gpu_codec_max_payload_bytes(0, 5) == 0


def test_product_never_imports_oracle():
    """The product package must not reach the oracle (no CPU fallback through the checker)."""
    pkg = os.path.join(ROOT, "image_compression_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".c", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "import oracle" not in text and "from oracle" not in text and "liboracle" not in text, f
