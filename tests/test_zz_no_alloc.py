"""GPU (-m gpu): gpu_codec_encode_bgr_dev_noalloc — the encode entry point for callers that
must not allocate device memory while encoding (their own pool owns the device, they cannot
take cudaMalloc's device-wide stall, or a capture is in progress elsewhere).  Its scratch is a
fixed array inside the library that every call on a device reuses, so the check is made on two
calls in a row: the second one sees the first one's leftovers and must still produce exactly
the bytes the regular entry point produces.  Collected last (test_zz_*) so a failure here
cannot hide the other tests under -x."""
import numpy as np
import pytest

from image_compression_b200 import synth

pytestmark = pytest.mark.gpu

KIND = {"noise": 0, "photo": 1, "flat": 2}


def _regular(codec, rows_t, w, h, stride):
    import torch
    payload = torch.empty(codec.max_payload_bytes(w, h) + 16, dtype=torch.uint8, device="cuda")
    n = torch.zeros(1, dtype=torch.int64, device="cuda")
    codec.encode_dev(0, rows_t, w, h, stride, payload, n)
    codec.sync_check(0)
    return payload[: int(n.item())].cpu().numpy()


def _noalloc(codec, rows_t, w, h, stride):
    import torch
    payload = torch.empty(codec.max_payload_bytes(w, h) + 16, dtype=torch.uint8, device="cuda")
    n = torch.full((1,), -1, dtype=torch.int64, device="cuda")
    status = torch.full((1,), 77, dtype=torch.int32, device="cuda")
    codec.encode_dev_noalloc(rows_t, w, h, stride, payload, n, status)
    torch.cuda.synchronize()
    assert int(status.item()) == 0
    return payload[: int(n.item())].cpu().numpy()


def test_noalloc_matches_regular_entry_point_twice_in_a_row(port):
    import torch
    from image_compression_b200 import codec
    torch.cuda.set_device(0)
    lib = codec.load()
    # call 1 (large, noise: many tiles, long chunks); the entry point uses no context (a fresh process that
    # never calls gpu_codec_init is exercised by test_noalloc_needs_no_init below)
    w1, h1 = 2048, 1024
    s1 = 3 * w1
    rows1 = torch.from_numpy(synth.frame_bgr_bottomup("noise", w1, h1, synth.SEED_BASE + 9).reshape(-1)).cuda()
    got1 = _noalloc(codec, rows1, w1, h1, s1)
    # call 2 straight after it (smaller, ragged, different content): stale descriptors of call 1 are in the array
    w2, h2 = 333, 77
    s2 = synth.bmp_stride(w2)
    rows2 = torch.from_numpy(synth.frame_bgr_bottomup("photo", w2, h2, synth.SEED_BASE + 9).reshape(-1)).cuda()
    got2 = _noalloc(codec, rows2, w2, h2, s2)
    # and a third, same input as the first again
    got3 = _noalloc(codec, rows1, w1, h1, s1)

    codec.init([0])
    want1 = _regular(codec, rows1, w1, h1, s1)
    want2 = _regular(codec, rows2, w2, h2, s2)
    assert got1.tobytes() == want1.tobytes()          # bit-exact, like every other parity test here
    assert got2.tobytes() == want2.tobytes()
    assert got3.tobytes() == want1.tobytes()
    # ... and both are the reference's bytes (the single-pass look-back placement against the oracle directly)
    assert got1.tobytes() == port.encode_bgr(rows1.cpu().numpy(), w1, h1, s1).tobytes()
    assert got2.tobytes() == port.encode_bgr(rows2.cpu().numpy(), w2, h2, s2).tobytes()

    # too small an output buffer is reported through d_status, nothing else breaks
    small = torch.empty(100, dtype=torch.uint8, device="cuda")
    n = torch.zeros(1, dtype=torch.int64, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    codec.encode_dev_noalloc(rows2, w2, h2, s2, small, n, status)
    torch.cuda.synchronize()
    assert int(status.item()) & 1

    # frames beyond the built-in scratch are refused up front with a clear message
    with pytest.raises(codec.GpuCodecError, match="built-in scratch"):
        codec.encode_dev_noalloc(rows2, 16 * 2048, 16 * 1025, 3 * 16 * 2048, small, n, status)
    assert b"2097152" in lib.gpu_codec_last_error()


def test_noalloc_needs_no_init():
    """A process that has never created a codec context (no gpu_codec_init) can still encode through this
    entry point: run in a child process of its own so that nothing else in the session has initialised the
    library first.  The child prints a digest of the payload, compared here with the regular entry point's."""
    import hashlib
    import os
    import subprocess
    import sys
    import torch
    from image_compression_b200 import codec
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    child = (
        "import hashlib, sys, torch\n"
        "sys.path.insert(0, %r)\n"
        "from image_compression_b200 import codec, synth\n"
        "torch.cuda.set_device(0)\n"
        "w, h = 640, 368; s = 3 * w\n"
        "rows = torch.from_numpy(synth.frame_bgr_bottomup('photo', w, h, synth.SEED_BASE + 9).reshape(-1)).cuda()\n"
        "out = []\n"
        "for _ in range(2):\n"
        "    payload = torch.empty(codec.max_payload_bytes(w, h) + 16, dtype=torch.uint8, device='cuda')\n"
        "    n = torch.zeros(1, dtype=torch.int64, device='cuda'); st = torch.ones(1, dtype=torch.int32, device='cuda')\n"
        "    codec.encode_dev_noalloc(rows, w, h, s, payload, n, st)\n"
        "    torch.cuda.synchronize()\n"
        "    assert int(st.item()) == 0 and codec.n_contexts() == 0\n"
        "    out.append(hashlib.sha256(payload[: int(n.item())].cpu().numpy().tobytes()).hexdigest())\n"
        "print('DIGESTS', *out)\n" % root)
    r = subprocess.run([sys.executable, "-c", child], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    digests = [ln for ln in r.stdout.splitlines() if ln.startswith("DIGESTS")][0].split()[1:]
    w, h = 640, 368
    s = 3 * w
    rows = torch.from_numpy(synth.frame_bgr_bottomup("photo", w, h, synth.SEED_BASE + 9).reshape(-1)).cuda()
    codec.init([0])
    want = hashlib.sha256(_regular(codec, rows, w, h, s).tobytes()).hexdigest()
    assert digests == [want, want]
