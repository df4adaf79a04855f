"""End-to-end encode from ordinary (pageable) host memory vs page-locked memory: gpu_codec_encode_bgr, 16384x16384."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from image_compression_b200 import codec, synth
codec.init([0])
lib = codec.load()
w = h = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
stride = 3 * w
n = stride * h
rows = torch.empty(n, dtype=torch.uint8, device="cuda")
codec.synth_dev(0, rows, w, h, stride, 1, synth.SEED_BASE + 3, 0)
codec.sync_check(0)
pageable = np.empty(n, np.uint8)
pageable[:] = rows.cpu().numpy()                       # touched, resident
hp = lib.gpu_codec_host_alloc(n)
pinned = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_uint8)), shape=(n,))
pinned[:] = pageable
res = {}
for name, buf in (("pinned", pinned), ("pageable", pageable)):
    out = None
    for i in range(5):
        t = time.perf_counter()
        addr, ln = codec.encode_bgr_ptr(buf.ctypes.data, w, h, stride)
        dt = time.perf_counter() - t
        got = bytes((C.c_uint8 * ln).from_address(addr)) if i == 0 else None
        codec.free_payload(addr)
        if i == 0:
            res[name] = got
        print(f"{name}: {dt*1e3:.2f} ms, {w*h/dt/1e9:.2f} Gpixel/s, payload {ln}")
print("identical payloads:", res["pinned"] == res["pageable"])

# decode: gpu_codec_decode_to_bgr into page-locked vs ordinary memory (payload from the same kind of memory)
pay = np.frombuffer(res["pinned"], np.uint8)
hq = lib.gpu_codec_host_alloc(pay.size)
pay_pinned = np.ctypeslib.as_array(C.cast(hq, C.POINTER(C.c_uint8)), shape=(pay.size,))
pay_pinned[:] = pay
pay_pageable = pay.copy()
out_pageable = np.zeros(n, np.uint8)
outs = {}
for name, p, o in (("pinned", pay_pinned, pinned), ("pageable", pay_pageable, out_pageable)):
    for i in range(4):
        t = time.perf_counter()
        codec.decode_to_bgr(p, w, h, stride, out=o)
        dt = time.perf_counter() - t
        print(f"decode {name}: {dt*1e3:.2f} ms, {w*h/dt/1e9:.2f} Gpixel/s")
    outs[name] = o.copy()
print("identical pixels:", bool((outs["pinned"] == outs["pageable"]).all()))
