import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def port():
    import oracle
    return oracle.port()


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference build; None where /root/reference (or a prebuilt oracle/_ref) is absent."""
    import oracle
    return oracle.ref()


@pytest.fixture(scope="session")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "reference_vectors.npz"))


@pytest.fixture(scope="session")
def emul():
    """Host build of csrc/codec_math.cuh (tests/emul): the kernels' arithmetic, checkable without a GPU."""
    import ctypes as C
    d = os.path.join(ROOT, "tests", "emul")
    so = os.path.join(d, "libemul.so")
    src = [os.path.join(d, "emul.cpp"), os.path.join(ROOT, "image_compression_b200", "csrc", "codec_math.cuh"),
           os.path.join(ROOT, "image_compression_b200", "csrc", "codec_tables.h")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
        subprocess.run(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-shared", "-fPIC", "-o", so, src[0]], check=True)
    lib = C.CDLL(so)
    lib.emul_transform_n.restype = C.c_long
    return lib


@pytest.fixture(scope="session")
def gpu():
    """Initialised libgpu_codec.so on cuda:0.  Fails (does not skip) when the extension or the GPU is missing."""
    from image_compression_b200 import codec
    codec.init([0])
    return codec


def golden_frames(golden):
    from image_compression_b200 import synth
    for i, m in enumerate(golden["meta"]):
        kind, w, h, seed, frame = str(m).split(",")
        w, h, seed, frame = int(w), int(h), int(seed), int(frame)
        yield i, kind, w, h, synth.frame_bgr_bottomup(kind, w, h, seed, frame)
