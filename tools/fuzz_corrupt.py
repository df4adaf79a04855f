"""Corruption campaign for the decoder (not part of the suite): random byte edits, truncations and extensions of valid
payloads.  The GPU decoder must never hang or fault; it must agree with the oracle: same pixels when the oracle
gets the expected chunks out of the bytes (whatever follows them is not read), an error otherwise.
usage: python tools/fuzz_corrupt.py [cases] [seed]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
from image_compression_b200 import codec, synth

n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 300
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 3)
port = oracle.port()
codec.init([0])
stats = {"both_ok": 0, "both_reject": 0, "disagree": 0, "pixel_mismatch": 0}
t0 = time.time()
for case in range(n_cases):
    w, h = int(rng.integers(16, 700)), int(rng.integers(16, 400))
    kind = ("photo", "noise")[int(rng.integers(0, 2))]
    stride = synth.bmp_stride(w)
    rows = synth.frame_bgr_bottomup(kind, w, h, 0xBAD0 + case)
    p = np.array(port.encode_bgr(rows, w, h, stride), np.uint8)
    op = int(rng.integers(0, 5))
    if op == 0:                                   # a few random byte values
        for _ in range(int(rng.integers(1, 4))):
            p[int(rng.integers(0, p.size))] = int(rng.integers(0, 256))
    elif op == 1:                                 # single bit flips
        for _ in range(int(rng.integers(1, 6))):
            p[int(rng.integers(0, p.size))] ^= 1 << int(rng.integers(0, 8))
    elif op == 2:                                 # truncate
        p = p[: int(rng.integers(0, p.size))].copy()
    elif op == 3:                                 # trailing bytes
        p = np.concatenate([p, rng.integers(0, 256, int(rng.integers(1, 40))).astype(np.uint8)])
    else:                                         # a run of zeros or 0xFF
        a = int(rng.integers(0, p.size)); n = int(rng.integers(1, 64))
        p[a: a + n] = 0 if rng.random() < 0.5 else 0xFF
    try:
        ref, used = port.decode(p, w, h, stride)
        oracle_ok = used <= p.size                # bytes after the last chunk are not read (awi.c:184-217)
    except ValueError:
        oracle_ok = False
    try:
        got = codec.decode_to_bgr(p, w, h, stride)
        gpu_ok = True
    except codec.GpuCodecError:
        gpu_ok = False
    if oracle_ok != gpu_ok:
        stats["disagree"] += 1
        print("DISAGREE", case, w, h, kind, op, "oracle", oracle_ok, "gpu", gpu_ok)
    elif gpu_ok:
        same = (got.reshape(h, stride)[:, : 3 * w] == ref.reshape(h, stride)[:, : 3 * w]).all()
        stats["both_ok" if same else "pixel_mismatch"] += 1
        if not same:
            print("PIXELS", case, w, h, kind, op)
    else:
        stats["both_reject"] += 1
# the context must still work
rows = synth.frame_bgr_bottomup("photo", 64, 48, 1)
ok = codec.encode_bgr(rows, 64, 48, synth.bmp_stride(64)) == port.encode_bgr(rows, 64, 48, synth.bmp_stride(64)).tobytes()
print(stats, "context still fine:", ok, f"{time.time() - t0:.0f} s")
sys.exit(0 if stats["disagree"] == 0 and stats["pixel_mismatch"] == 0 and ok else 1)
