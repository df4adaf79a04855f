"""Randomised parity campaign (not part of the test suite): encode bytes, chunk offsets and decoded pixels of the
GPU path against the CPU oracle over random geometry / content / stride / entry point, plus the sizes where the
code changes strategy (bulk-copy tiles, 4096-tile scan+gather boundary, 1024-tile DC boundary).
usage: python tools/fuzz_parity.py [cases] [seed]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import oracle
from image_compression_b200 import codec, synth

n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 200
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
port = oracle.port()
codec.init([0])
t0 = time.time()
fixed = [(8192, 4096 - 16), (8192, 4096), (8192, 4096 + 16), (4096, 4096), (4096 + 16, 4096), (16 * 512, 16 * 257),
         (512, 16), (512 + 16, 16), (16 * 33, 16 * 33), (16 * 1024 + 7, 40)]
bad = 0
for case in range(n_cases + len(fixed)):
    if case < len(fixed):
        w, h = fixed[case]
    else:
        big = rng.random() < 0.15
        w = int(rng.integers(1, 2600 if not big else 6000)); h = int(rng.integers(1, 1400 if not big else 2500))
    kind = ("photo", "noise", "flat")[int(rng.integers(0, 3))] if case >= len(fixed) else "photo"
    stride = synth.bmp_stride(w) + (0 if rng.random() < 0.6 else int(rng.integers(0, 5)) * 4 + int(rng.integers(0, 2)) * int(rng.integers(0, 4)))
    base = synth.frame_bgr_bottomup(kind, w, h, 0xF00D + case)
    rows = np.zeros(stride * h, np.uint8)
    rows.reshape(h, stride)[:, : synth.bmp_stride(w)] = base.reshape(h, synth.bmp_stride(w))
    want, _, lens = port.encode_bgr(rows, w, h, stride, want_coefs=True)
    want = np.asarray(want, np.uint8)
    mode = int(rng.integers(0, 3))
    if mode == 0:
        got = np.frombuffer(codec.encode_bgr(rows, w, h, stride), np.uint8)
    else:
        rt = torch.from_numpy(rows).cuda()
        nch = lens.size
        payload = torch.empty(nch * 121 + 16, dtype=torch.uint8, device="cuda")
        ln = torch.zeros(1, dtype=torch.int64, device="cuda")
        off = torch.empty(nch, dtype=torch.int32, device="cuda")
        if mode == 1:
            codec.encode_dev(0, rt, w, h, stride, payload, ln, chunk_off=off)
        else:
            st = torch.zeros(1, dtype=torch.int32, device="cuda")
            codec.encode_dev_noalloc(rt, w, h, stride, payload, ln, st)
        codec.sync_check(0)
        got = payload[: int(ln.item())].cpu().numpy()
        if mode == 1 and not (off.cpu().numpy().view(np.uint32) == np.concatenate([[0], np.cumsum(lens)[:-1]])).all():
            print("OFFSET MISMATCH", case, w, h, kind); bad += 1
    if got.tobytes() != want.tobytes():
        print("ENCODE MISMATCH", case, w, h, kind, stride, mode); bad += 1
        continue
    ref_px = port.decode(want, w, h, stride)[0]
    px = codec.decode_to_bgr(want, w, h, stride)
    if not (px.reshape(h, stride)[:, : 3 * w] == ref_px.reshape(h, stride)[:, : 3 * w]).all():
        print("DECODE MISMATCH", case, w, h, kind, stride); bad += 1
    if case % 25 == 0:
        print(f"case {case}: {w}x{h} {kind} ok ({time.time() - t0:.0f} s)", flush=True)
print("mismatches:", bad, "cases:", n_cases + len(fixed), f"{time.time() - t0:.0f} s")
sys.exit(1 if bad else 0)
