"""Measures pinned host<->device copy bandwidth on the GPU box (context for the e2e number)."""
import time, torch
n = 805306368
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
for name, fn in (("H2D", lambda: d.copy_(h, non_blocking=True)), ("D2H", lambda: h.copy_(d, non_blocking=True))):
    fn(); torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(5): fn()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t) / 5
    print(f"{name} pinned {n/dt/1e9:.1f} GB/s ({dt*1e3:.1f} ms for {n/1e6:.0f} MB)")
p = torch.empty(n, dtype=torch.uint8)
t = time.perf_counter(); d.copy_(p); torch.cuda.synchronize(); dt = time.perf_counter() - t
print(f"H2D pageable {n/dt/1e9:.1f} GB/s")
import os; print("cpus", os.cpu_count())
