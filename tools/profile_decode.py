"""Small driver for profiling the decode kernels: 8192x8192 (configs[4]) encode once, decode N times."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from image_compression_b200 import codec, synth
kind = sys.argv[1] if len(sys.argv) > 1 else "photo"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
codec.init([0])
w = h = 8192
stride = 3 * w
rows = torch.empty(h * stride, dtype=torch.uint8, device="cuda")
codec.synth_dev(0, rows, w, h, stride, {"noise": 0, "photo": 1}[kind], synth.SEED_BASE + 5, 0)
payload = torch.empty(codec.max_payload_bytes(w, h) + 16, dtype=torch.uint8, device="cuda")
plen = torch.zeros(1, dtype=torch.int64, device="cuda")
codec.encode_dev(0, rows, w, h, stride, payload, plen)
codec.sync_check(0)
nbytes = int(plen.item())
dst = torch.empty(h * stride, dtype=torch.uint8, device="cuda")
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(n):
    a.record()
    codec.decode_dev(0, payload, nbytes, w, h, dst, stride)
    b.record()
    codec.sync_check(0)
    print(f"decode {kind} {w}x{h}: {a.elapsed_time(b):.3f} ms, payload {nbytes} bytes")
