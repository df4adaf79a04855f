"""How much pinned host->device bandwidth do k GPUs of one box get when they all upload at once?
(Context for the end-to-end number at N > 1: every rank uploads 805 MB per frame.)
Run under torchrun with one rank per GPU:  torchrun --nproc-per-node 8 tools/pcie_concurrency.py"""
import os, time
import torch
import torch.distributed as dist

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
n = 805306368
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
d.copy_(h, non_blocking=True); torch.cuda.synchronize()
flag = torch.zeros(1, device="cuda")
k = 1
while k <= world:
    if world > 1:
        dist.all_reduce(flag)                       # barrier on the device
    torch.cuda.synchronize()
    gbs = 0.0
    if rank < k:
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(8):
            d.copy_(h, non_blocking=True)
        b.record(); torch.cuda.synchronize()
        gbs = 8 * n / (a.elapsed_time(b) * 1e-3) / 1e9
    t = torch.tensor([gbs], device="cuda")
    if world > 1:
        dist.all_reduce(t)
    if rank == 0:
        print(f"{k} GPU(s) uploading at once: {t.item():.1f} GB/s aggregate, {t.item() / k:.1f} per GPU", flush=True)
    k *= 2
if rank == 0:
    print("cpus", os.cpu_count())
if world > 1:
    dist.destroy_process_group()
