"""Small driver for profiling the encode kernels: 16384x16384 (configs[2]) resident in HBM, encode N times.
  python tools/profile_encode.py [photo|noise] [N] [width height]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from image_compression_b200 import codec, synth
kind = sys.argv[1] if len(sys.argv) > 1 else "photo"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
w = int(sys.argv[3]) if len(sys.argv) > 4 else 16384
h = int(sys.argv[4]) if len(sys.argv) > 4 else 16384
codec.init([0])
stride = 3 * w
rows = torch.empty(h * stride, dtype=torch.uint8, device="cuda")
codec.synth_dev(0, rows, w, h, stride, {"noise": 0, "photo": 1}[kind], synth.SEED_BASE + 3, 0)
payload = torch.empty(codec.max_payload_bytes(w, h) + 16, dtype=torch.uint8, device="cuda")
plen = torch.zeros(1, dtype=torch.int64, device="cuda")
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(n):
    a.record()
    codec.encode_dev(0, rows, w, h, stride, payload, plen)
    b.record()
    codec.sync_check(0)
    print(f"encode {kind} {w}x{h}: {a.elapsed_time(b):.4f} ms, payload {int(plen.item())} bytes")
