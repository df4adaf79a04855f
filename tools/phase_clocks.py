"""Per-warp cycles between the phase marks of awi_encode_kernel (16384x16384, two-pass mode).

Needs a library built with the marks compiled in:
    AWI_NVCC_EXTRA=-DENC_PHASE_CLOCKS python image_compression_b200/build.py --force
(rebuild without the flag afterwards; the marks cost time and the shipped library has none).
"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from image_compression_b200 import codec, synth

kind = sys.argv[1] if len(sys.argv) > 1 else "photo"
codec.init([0])
lib = ctypes.CDLL(codec.LIB_PATH if hasattr(codec, "LIB_PATH") else os.path.join(os.path.dirname(codec.__file__), "libgpu_codec.so"))
w = h = 16384
stride = 3 * w
rows = torch.empty(h * stride, dtype=torch.uint8, device="cuda")
codec.synth_dev(0, rows, w, h, stride, {"noise": 0, "photo": 1}[kind], synth.SEED_BASE + 3, 0)
payload = torch.empty(codec.max_payload_bytes(w, h) + 16, dtype=torch.uint8, device="cuda")
plen = torch.zeros(1, dtype=torch.int64, device="cuda")
for _ in range(2):
    codec.encode_dev(0, rows, w, h, stride, payload, plen)
codec.sync_check(0)
buf = (ctypes.c_ulonglong * 128)()
assert lib.awi_debug_phase_clocks(None, 1)
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
codec.encode_dev(0, rows, w, h, stride, payload, plen)
b.record()
codec.sync_check(0)
assert lib.awi_debug_phase_clocks(buf, 0)
n_tiles = (w // 16) * (h // 16) // 32
names = ["A stage", "wait A->B", "B colour+dct+quant", "halo", "wait B->C", "C exact", "wait C->D", "D rle", "E warp scan",
         "wait E", "compact copy", "wait compact", "store"]
print(f"{kind}: {a.elapsed_time(b):.3f} ms with marks; kcycles per tile per warp (warps 0-5), {n_tiles} tiles")
tot = [0.0] * 6
for i, nm in enumerate(names):
    v = [buf[i * 8 + wp] / n_tiles / 1e3 for wp in range(6)]
    tot = [t + x for t, x in zip(tot, v)]
    print(f"{nm:22s}" + "".join(f"{x:8.2f}" for x in v))
print(f"{'sum':22s}" + "".join(f"{x:8.2f}" for x in tot))
